"""transformer-stm_b200: B200-native CvT hot path.  Import as `transformer_stm_b200`."""
from ._lib import CvtError, LIB_PATH, exported_names  # noqa: F401
from .engine import CvtConfig, CvtEngine, StageSpec  # noqa: F401
