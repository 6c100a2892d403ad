from .mixup import MetaMixup, AttnMixup, mixup_data, mixup_criterion, mixup_criterion_meta, sample_beta  # noqa: F401
from .synthetic import SyntheticCIFAR, SyntheticAGNews  # noqa: F401
from .loader import BackgroundLoader, TensorBatchLoader, DevicePrefetcher  # noqa: F401
