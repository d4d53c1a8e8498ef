from .resnet import NetResDeep, ResBlock  # noqa: F401
