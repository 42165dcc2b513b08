#!/usr/bin/env python
"""MNIST CNN, session-style loop — the b200dp counterpart of the script the MPIJob
runs (reference ``horovod/tensorflow_mnist.py``; same flags: ``--use-adasum --lr --num-steps``).

    b200run -np 8 python examples/tensorflow_mnist.py --num-steps 20000
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import b200dp as hvd
from b200dp.data import mnist

parser = argparse.ArgumentParser(description="b200dp MNIST example (session loop)")
parser.add_argument("--use-adasum", action="store_true", default=False, help="use adasum algorithm to do reduction")
parser.add_argument("--lr", default=0.001, type=float, help="Adam learning rate")
parser.add_argument("--num-steps", default=20000, type=int, help="Number of training steps")
# extensions (reference hard-codes these)
parser.add_argument("--batch-size", default=100, type=int, help="per-rank batch (reference: 100)")
parser.add_argument("--arch", default="mnist_cnn", choices=["mnist_cnn", "keras128"])
parser.add_argument("--width", default=1.0, type=float, help="channel/dense width multiplier")
parser.add_argument("--synthetic", action="store_true", help="synthetic 28x28 data (no dataset file needed)")
parser.add_argument("--checkpoint-dir", default="./checkpoints")
parser.add_argument("--checkpoint-secs", default=600.0, type=float)
parser.add_argument("--log-every", default=10, type=int)
parser.add_argument("--engine", default="auto", choices=["auto", "b200", "torch"])
parser.add_argument("--allreduce-algo", default="auto", choices=["auto", "oneshot", "twoshot", "nvls"])
parser.add_argument("--train-samples", default=60000, type=int)


def main():
    args = parser.parse_args()
    hvd.init()

    (x_train, y_train), _ = mnist.load_data("MNIST-data-%d" % hvd.rank(), synthetic=args.synthetic or None,
                                            n_train=args.train_samples, n_test=1000)
    x_train = mnist.preprocess(x_train)

    lr_scaler = hvd.lr_scaler(args.use_adasum)
    opt = hvd.Adam(lr=args.lr * lr_scaler)
    opt = hvd.DistributedOptimizer(opt, op=hvd.Adasum if args.use_adasum else hvd.Average, algo=args.allreduce_algo)

    engine = hvd.make_engine(hvd.get_spec(args.arch, args.width), opt, engine=args.engine,
                             batch_size=args.batch_size)
    hooks = [
        hvd.BroadcastGlobalVariablesHook(0),
        hvd.StopAtStepHook(last_step=args.num_steps // hvd.size()),
        hvd.LoggingTensorHook(tensors=("step", "loss"), every_n_iter=args.log_every),
    ]
    checkpoint_dir = args.checkpoint_dir if hvd.rank() == 0 else None
    batches = mnist.train_input_generator(x_train, y_train, batch_size=args.batch_size)
    with hvd.MonitoredTrainingSession(engine, checkpoint_dir=checkpoint_dir, hooks=hooks,
                                      save_checkpoint_secs=args.checkpoint_secs) as sess:
        while not sess.should_stop():
            image_, label_ = next(batches)
            sess.run(image_, label_)
    hvd.shutdown()


if __name__ == "__main__":
    main()
