"""b200dp — Blackwell-native data-parallel training with a Horovod-shaped API.

``import b200dp as hvd`` gives the names the reference scripts use
(``horovod/tensorflow_mnist.py``, ``horovod/tensorflow_mnist_gpu.py``):
``init rank size local_rank local_size nccl_built Average Adasum
DistributedOptimizer BroadcastGlobalVariablesHook callbacks.*`` plus the raw
``allreduce`` / ``broadcast`` primitives.
"""
from .runtime.world import (init, shutdown, is_initialized, rank, size, local_rank, local_size,
                            nccl_built, mpi_built, gloo_built, barrier, device, world)
from .parallel.collectives import (ReduceOp, Average, Sum, Adasum, allreduce, allreduce_, broadcast,
                                   broadcast_, allgather, broadcast_object, allgather_object,
                                   broadcast_parameters, broadcast_optimizer_state)
from .parallel.collectives import Min, Max, Product
from .parallel.extras import (Handle, poll, synchronize, allreduce_async, allreduce_async_, broadcast_async, broadcast_async_,
                              allgather_async, grouped_allreduce, grouped_allreduce_, alltoall, reducescatter, join,
                              Compression, mpi_threads_supported, mpi_enabled, gloo_enabled, cuda_built, rocm_built,
                              ccl_built, ddl_built, cross_rank, cross_size)
from .parallel.optimizer import Adam, DistributedAdam, DistributedOptimizer, lr_scaler
from .train.session import (MonitoredTrainingSession, SessionRunHook, BroadcastGlobalVariablesHook,
                            StopAtStepHook, LoggingTensorHook, StepCounterHook, CheckpointSaverHook, MetricsHook)
from .train.trainer import Model, MNISTModel, SparseCategoricalCrossentropy
from .train.engine import make_engine, TorchEngine, StepResult
from .models.spec import get_spec, ModelSpec, LayerSpec, custom as custom_spec
from . import callbacks
from . import elastic

__version__ = "0.1.0"
