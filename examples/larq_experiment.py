"""Zookeeper-configured Larq BinaryNet experiment, running on the B200-native engine.

Counterpart of the reference's examples/larq_experiment.py: same components, fields, task name and
CLI (``python examples/larq_experiment.py BinaryNetMnist epochs=1 batch_size=4096``); the TF / Larq
calls are replaced by the symbolic ops and layer descriptions of ``zookeeper_b200.tf``.
"""
import os
import sys
from functools import partial
from typing import Sequence, Tuple, Union

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from zookeeper_b200 import ComponentField, Field, cli, component, factory, task  # noqa: E402
from zookeeper_b200.tf import Dataset, Experiment, Preprocessing, SyntheticDataset, TFDSDataset  # noqa: E402
from zookeeper_b200.tf import image as zimage  # noqa: E402
from zookeeper_b200.tf import layers as lq_layers  # noqa: E402
from zookeeper_b200.tf.models import Model, Sequential  # noqa: E402
from zookeeper_b200.tf.optimizers import Adam  # noqa: E402


@component
class Mnist(TFDSDataset):
    """MNIST from ``<data_dir>/mnist.npz`` (keys x_train, y_train, x_test, y_test)."""

    name = Field("mnist")
    train_split = Field("train")
    validation_split = Field("test")


@component
class SyntheticMnist(SyntheticDataset):
    """MNIST-shaped synthetic data (no dataset files are needed)."""

    image_shape = Field((28, 28, 1))
    num_classes = Field(10)


@component
class PadCropAndFlip(Preprocessing):
    pad_size: int = Field()

    def input(self, data, training):
        image = data["image"]
        if training:
            image = zimage.resize_with_crop_or_pad(image, self.pad_size, self.pad_size)
            image = zimage.random_crop(image, self.input_shape)
            image = zimage.random_flip_left_right(image)
        else:
            image = zimage.resize_with_crop_or_pad(image, *self.input_shape[:2])
        return zimage.cast(image, "float32") / (255.0 / 2.0) - 1.0

    def output(self, data):
        return data["label"]


@factory
class BinaryNet:
    dataset: Dataset = ComponentField()
    preprocessing: Preprocessing = ComponentField()

    filters: int = Field(128)
    dense_units: int = Field(1024)
    kernel_size: Union[int, Tuple[int, int]] = Field((3, 3))

    input_shape: Tuple[int, int, int] = Field()

    def build(self) -> Model:
        kwhparams = dict(
            input_quantizer="ste_sign",
            kernel_quantizer="ste_sign",
            kernel_constraint="weight_clip",
            use_bias=False,
        )
        L = lq_layers
        return Sequential(
            [
                # Don't quantize inputs in first layer
                L.QuantConv2D(
                    self.filters,
                    self.kernel_size,
                    kernel_quantizer="ste_sign",
                    kernel_constraint="weight_clip",
                    use_bias=False,
                    input_shape=self.input_shape,
                ),
                L.BatchNormalization(scale=False),
                L.QuantConv2D(self.filters, self.kernel_size, padding="same", **kwhparams),
                L.MaxPool2D(pool_size=(2, 2), strides=(2, 2)),
                L.BatchNormalization(scale=False),
                L.QuantConv2D(2 * self.filters, self.kernel_size, padding="same", **kwhparams),
                L.BatchNormalization(scale=False),
                L.QuantConv2D(2 * self.filters, self.kernel_size, padding="same", **kwhparams),
                L.MaxPool2D(pool_size=(2, 2), strides=(2, 2)),
                L.BatchNormalization(scale=False),
                L.QuantConv2D(4 * self.filters, self.kernel_size, padding="same", **kwhparams),
                L.BatchNormalization(scale=False),
                L.QuantConv2D(4 * self.filters, self.kernel_size, padding="same", **kwhparams),
                L.MaxPool2D(pool_size=(2, 2), strides=(2, 2)),
                L.BatchNormalization(scale=False),
                L.Flatten(),
                L.QuantDense(self.dense_units, **kwhparams),
                L.BatchNormalization(scale=False),
                L.QuantDense(self.dense_units, **kwhparams),
                L.BatchNormalization(scale=False),
                L.QuantDense(self.dataset.num_classes, **kwhparams),
                L.BatchNormalization(scale=False),
                L.Activation("softmax"),
            ]
        )


@task
class BinaryNetMnist(Experiment):
    dataset = ComponentField(SyntheticMnist)
    input_shape: Tuple[int, int, int] = Field((28, 28, 1))
    preprocessing = ComponentField(PadCropAndFlip, pad_size=32)
    model: Model = ComponentField(BinaryNet)

    epochs = Field(100)
    batch_size = Field(128)
    learning_rate: float = Field(5e-3)

    loss = Field("sparse_categorical_crossentropy")
    metrics: Sequence[str] = Field(lambda: ["accuracy"])

    @Field
    def optimizer(self):
        return Adam(self.learning_rate)

    def run(self):
        _maybe_init_distributed()
        train_data, num_train_examples = self.dataset.train()
        train_data = (
            train_data.cache()
            .shuffle(10 * self.batch_size)
            .repeat()
            .map(partial(self.preprocessing, training=True))
            .batch(self.batch_size)
        )
        validation_data, num_validation_examples = self.dataset.validation()
        validation_data = (
            validation_data.cache()
            .repeat()
            .map(self.preprocessing)
            .batch(self.batch_size)
        )

        self.model.compile(optimizer=self.optimizer, loss=self.loss, metrics=self.metrics)

        self.model.summary()

        self.model.fit(
            train_data,
            epochs=self.epochs,
            steps_per_epoch=num_train_examples // self.batch_size,
            validation_data=validation_data,
            validation_steps=num_validation_examples // self.batch_size,
        )


def _maybe_init_distributed():
    """One process per GPU when launched with torch.distributed.run."""
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch
        import torch.distributed as dist

        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))


if __name__ == "__main__":
    cli()
