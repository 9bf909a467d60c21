"""Stand-in so zipvoice/utils/_bigvgan_mel_feature.py imports; BigVGAN features are out of scope."""
