"""atlas_vae_b200: B200-native OE-VAE hot path (training step, anomaly scoring, cut + mass histogram).

Drop-in for the reference's OE-VAE/models.py + the hot-path parts of utils.py / plots.py / vae.py, backed by
liboevae.so (hand-written sm_100a CUDA behind the C ABI of include/oevae.h).  There is no CPU fallback.
"""
from .models import VariationalAutoEncoder, train_model, get_losses, model_checkpoint  # noqa: F401
