def mel(*a, **k):
    raise NotImplementedError("librosa is not available; BigVGAN features are out of scope")
