class FeatureExtractor:
    """lhotse.features.base.FeatureExtractor: stores ``config`` (feature.py:44)."""
    name = None
    config_type = None

    def __init__(self, config=None):
        self.config = config if config is not None else self.config_type()


def register_extractor(cls):
    return cls
