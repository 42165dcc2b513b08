#!/usr/bin/env python
"""MNIST CNN, Keras-style fit with bf16 mixed precision — the b200dp counterpart of
reference ``horovod/tensorflow_mnist_gpu.py`` (flags ``--use-adasum --lr --num-steps --batch-size``).

    b200run -np 8 python examples/tensorflow_mnist_gpu.py --num-steps 2000 --batch-size 128
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import b200dp as hvd
from b200dp.data import mnist

parser = argparse.ArgumentParser(description="b200dp MNIST example (Keras-style fit)")
parser.add_argument("--use-adasum", action="store_true", default=False, help="use adasum algorithm to do reduction")
parser.add_argument("--lr", default=0.001, type=float, help="Adam learning rate")
parser.add_argument("--num-steps", default=20000, type=int, help="Number of training steps")
parser.add_argument("--batch-size", default=100, type=int, help="Batch size")
parser.add_argument("--arch", default="mnist_cnn", choices=["mnist_cnn", "keras128"])
parser.add_argument("--width", default=1.0, type=float)
parser.add_argument("--synthetic", action="store_true")
parser.add_argument("--engine", default="auto", choices=["auto", "b200", "torch"])
parser.add_argument("--allreduce-algo", default="auto", choices=["auto", "oneshot", "twoshot", "nvls"])
parser.add_argument("--log-dir", default="./logs")
parser.add_argument("--checkpoint-dir", default="./checkpoints")
parser.add_argument("--export-dir", default="./final_model")
parser.add_argument("--train-samples", default=60000, type=int)
parser.add_argument("--test-samples", default=10000, type=int)
parser.add_argument("--resume", action="store_true",
                    help="continue from the newest epoch checkpoint in --checkpoint-dir (implied when b200run restarts the job)")


def main():
    args = parser.parse_args()
    hvd.init()          # also pins this process to GPU local_rank (one GPU per process)

    (x_train, y_train), (x_test, y_test) = mnist.load_data(
        f"MNIST-data-{hvd.rank()}", synthetic=args.synthetic or None,
        n_train=args.train_samples, n_test=args.test_samples)
    x_train, x_test = mnist.preprocess(x_train), mnist.preprocess(x_test)

    model = hvd.MNISTModel(width=args.width, arch=args.arch, engine=args.engine, dtype="bf16",
                           batch_size=args.batch_size)

    lr_scaler = hvd.lr_scaler(args.use_adasum)
    opt = hvd.Adam.keras(lr=args.lr * lr_scaler)
    opt = hvd.DistributedOptimizer(opt, op=hvd.Adasum if args.use_adasum else hvd.Average, algo=args.allreduce_algo)
    model.compile(optimizer=opt, loss=hvd.SparseCategoricalCrossentropy(from_logits=True), metrics=["accuracy"])

    callbacks = [
        hvd.callbacks.BroadcastGlobalVariablesCallback(0),
        hvd.callbacks.MetricAverageCallback(),
    ]
    if hvd.rank() == 0:
        callbacks.append(hvd.callbacks.TensorBoard(log_dir=args.log_dir))
        callbacks.append(hvd.callbacks.ModelCheckpoint(
            os.path.join(args.checkpoint_dir, "mnist-{epoch}.safetensors"), save_best_only=True))

    # The reference script has no resume path; `b200run --max-restarts` relaunches the job after a rank died, and a relaunch
    # that began at epoch 0 again would quietly throw the finished epochs away.  Rank 0 reloads the newest epoch checkpoint
    # (weights, Adam moments, step); the broadcast callback hands it to everyone on the first batch.
    initial_epoch = 0
    if args.resume or int(os.environ.get("B200DP_RESTART_COUNT", "0")) > 0:
        if hvd.rank() == 0 and os.path.isdir(args.checkpoint_dir):
            import re
            found = sorted((int(m.group(1)), f) for f in os.listdir(args.checkpoint_dir)
                           for m in [re.match(r"mnist-(\d+)\.safetensors$", f)] if m)
            if found:
                initial_epoch, name = found[-1]
                model.load_weights(os.path.join(args.checkpoint_dir, name))
                print(f"resuming from {name} (epoch {initial_epoch})", file=sys.stderr)
        initial_epoch = hvd.broadcast_object(initial_epoch, 0)

    steps_per_epoch = max(1, len(x_train) // (args.batch_size * hvd.size()))
    validation_steps = max(1, len(x_test) // args.batch_size)
    epochs = max(1, args.num_steps // steps_per_epoch)

    model.fit(x_train, y_train, batch_size=args.batch_size, epochs=epochs, steps_per_epoch=steps_per_epoch,
              validation_data=(x_test, y_test), validation_steps=validation_steps, callbacks=callbacks,
              verbose=1 if hvd.rank() == 0 else 0, initial_epoch=initial_epoch)

    if hvd.rank() == 0:
        score = model.evaluate(x_test, y_test, batch_size=args.batch_size, verbose=0)
        print("Test loss:", score[0])
        print("Test accuracy:", score[1])
        model.save(args.export_dir)
    hvd.shutdown()


if __name__ == "__main__":
    main()
