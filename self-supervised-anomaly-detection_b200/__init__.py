"""B200-native artificial-anomaly generator (drop-in for the data path of
gabry1998/Self-Supervised-Anomaly-Detection: ``src/self_supervised/datasets.py`` +
``dataset_generator.py``).  Importable as ``ssad_b200`` through the root shim."""
from .config import CPP, GeneratorConfig  # noqa: F401
from .params import ParamBatch, sample_params, params_from_draws, IMAGE_PARAMS_DTYPE  # noqa: F401
from . import constants  # noqa: F401

__version__ = "0.1.0"
