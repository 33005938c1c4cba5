"""b200sd — B200-native Stable-Diffusion-class batch text-to-image hot path.

Drop-in for the compute behind christianshub/k8s-nvidia-gpus' sd15-api (`pipe(...)` in
cluster-config/apps/sd15-api/configmap.yaml:105-112) and its scripts/batch_generate.py client.
The directory name contains '-' so import it with importlib (see `b200sd.py` at the repo root).
"""
from . import lib  # noqa: F401
from . import ops  # noqa: F401
from . import tokenizer  # noqa: F401
from . import weights  # noqa: F401
from . import checkpoint  # noqa: F401
from . import schedulers  # noqa: F401
from . import engine  # noqa: F401
from . import clip  # noqa: F401
from . import unet  # noqa: F401
from . import vae  # noqa: F401
from . import pipeline  # noqa: F401
from . import dist  # noqa: F401
from .pipeline import StableDiffusionPipeline, StableDiffusionXLPipeline  # noqa: F401

__version__ = "0.1.0"
