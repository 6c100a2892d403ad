from .ngd import NGD, FactorBatch  # noqa: F401
from .madgrad import MADGRAD, MirrorMADGRAD  # noqa: F401
from .schedules import build_scheduler  # noqa: F401
