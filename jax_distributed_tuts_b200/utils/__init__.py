from .metrics import get_num_params, print_exception, print_metrics  # noqa: F401
from .profiling import named_scope  # noqa: F401
