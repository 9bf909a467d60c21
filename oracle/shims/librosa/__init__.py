"""Stand-in for ``librosa`` (third-party, absent from this image) so that the unmodified
zipvoice/utils/_bigvgan_mel_feature.py imports: only ``librosa.filters.mel`` is provided (filters.py)."""
from . import filters  # noqa: F401
