"""Float64 export twin loaded by save_model_to_torch_script.py (see export_double.py)."""
from .export_double import AC_IQN_Actor as Actor  # noqa: F401
