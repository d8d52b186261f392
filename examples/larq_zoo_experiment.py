"""Zookeeper-configured experiments on the two larq-zoo models BASELINE.json names besides BinaryNet
(configs[3] BinaryAlexNet, configs[4] Bi-Real-Net-18), running on the B200-native engines.

The shape of the reference's examples/larq_experiment.py -- dataset / preprocessing / model
components, a ``@task`` experiment whose ``run()`` builds the tf.data-style pipeline, compiles and fits
-- with the model factories returning the zoo models instead of the example's BinaryNet:

    python examples/larq_zoo_experiment.py BinaryAlexNetImagenet epochs=1 batch_size=256
    python examples/larq_zoo_experiment.py BiRealNet18Imagenet epochs=1 batch_size=256
    python -m torch.distributed.run --nproc-per-node 8 examples/larq_zoo_experiment.py BinaryAlexNetImagenet batch_size=2048

Training only (the zoo engines have no inference path yet, so ``run()`` fits without validation
data); ``epochs=0`` configures, describes and lowers everything without a GPU.
"""
import os
import sys
from functools import partial
from typing import Sequence, Tuple

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from zookeeper_b200 import ComponentField, Field, cli, component, factory, task  # noqa: E402
from zookeeper_b200.tf import Dataset, Experiment, Preprocessing, SyntheticDataset  # noqa: E402
from zookeeper_b200.tf import image as zimage  # noqa: E402
from zookeeper_b200.tf import zoo  # noqa: E402
from zookeeper_b200.tf.models import Model  # noqa: E402
from zookeeper_b200.tf.optimizers import Adam  # noqa: E402


@component
class SyntheticImagenet(SyntheticDataset):
    """ImageNet-shaped synthetic data: 256x256x3 uint8 images (the usual resize before the 224 crop),
    1000 classes.  2048 images are 400 MB of pool; no dataset files are needed."""

    image_shape = Field((256, 256, 3))
    num_classes = Field(1000)
    num_train_examples = Field(2048)
    num_validation_examples = Field(0)


@component
class RandomCropAndFlip(Preprocessing):
    """Training: random ``input_shape`` crop of the stored image + random left-right flip; otherwise
    the centre crop; then ``x / 127.5 - 1``."""

    def input(self, data, training):
        image = data["image"]
        if training:
            image = zimage.random_crop(image, self.input_shape)
            image = zimage.random_flip_left_right(image)
        else:
            image = zimage.resize_with_crop_or_pad(image, *self.input_shape[:2])
        return zimage.cast(image, "float32") / (255.0 / 2.0) - 1.0

    def output(self, data):
        return data["label"]


@factory
class BinaryAlexNet:
    dataset: Dataset = ComponentField()
    input_shape: Tuple[int, int, int] = Field()
    dense_units: int = Field(4096)

    def build(self) -> Model:
        return zoo.BinaryAlexNet(self.input_shape, self.dataset.num_classes, self.dense_units)


@factory
class BiRealNet18:
    dataset: Dataset = ComponentField()
    input_shape: Tuple[int, int, int] = Field()

    def build(self) -> Model:
        return zoo.BiRealNet18(self.input_shape, self.dataset.num_classes)


class _ZooExperiment(Experiment):
    dataset = ComponentField(SyntheticImagenet)
    input_shape: Tuple[int, int, int] = Field((224, 224, 3))
    preprocessing = ComponentField(RandomCropAndFlip)

    epochs = Field(1)
    batch_size = Field(256)
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
        self.model.compile(optimizer=self.optimizer, loss=self.loss, metrics=self.metrics)
        self.model.summary()
        self.model.fit(train_data, epochs=self.epochs, steps_per_epoch=num_train_examples // self.batch_size)


@task
class BinaryAlexNetImagenet(_ZooExperiment):
    model: Model = ComponentField(BinaryAlexNet)


@task
class BiRealNet18Imagenet(_ZooExperiment):
    model: Model = ComponentField(BiRealNet18)


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
