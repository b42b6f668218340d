"""b200dp — Blackwell-native data-parallel training framework.

Capabilities mirror the reference project (three training modes, MobileNetV2 transfer
learning with a 10-class head, CrossEntropy + Adam + StepLR, partitioned sampler,
rank-0 metric collection, torchvision-format checkpoints, top-k / web inference), but the
architecture is B200-first: one process per GPU, flat parameter arenas, hand-written
sm_100a kernels behind a C ABI (``include/b200dp.h``), a fused peer-memory
all-reduce + Adam step, and CUDA-event timing.

Import as ``b200dp`` (see ``b200dp/__init__.py`` at the repository root).
"""

__version__ = "0.1.0"

CIFAR10_CLASSES = (
    "airplane", "automobile", "bird", "cat", "deer",
    "dog", "frog", "horse", "ship", "truck",
)  # reference: cifar10_mpi_mobilenet_224.py:65-66

IMAGENET_MEAN = (0.485, 0.456, 0.406)  # reference: cifar10_mpi_mobilenet_224.py:80
IMAGENET_STD = (0.229, 0.224, 0.225)   # reference: cifar10_mpi_mobilenet_224.py:81
