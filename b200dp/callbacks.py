"""``hvd.callbacks.*`` plus the Keras callbacks the reference attaches on rank 0
(``horovod/tensorflow_mnist_gpu.py:147-163``)."""
from .train.trainer import (Callback, BroadcastGlobalVariablesCallback, MetricAverageCallback,
                            TensorBoard, ModelCheckpoint, History, MetricsCallback,
                            LearningRateScheduleCallback, LearningRateWarmupCallback)

__all__ = ["Callback", "BroadcastGlobalVariablesCallback", "MetricAverageCallback",
           "TensorBoard", "ModelCheckpoint", "History", "MetricsCallback",
           "LearningRateScheduleCallback", "LearningRateWarmupCallback"]
