from .trainer import Trainer, as_engine  # noqa: F401
from .metrics import DeviceMetrics, StepTimer, draw_graph, peak_memory_gb  # noqa: F401
from .checkpoint import save_checkpoint, load_best_performance, load_weights  # noqa: F401
