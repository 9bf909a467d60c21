"""zipvoice_b200: B200-native (sm_100a) implementation of the ZipVoice inference hot path.

Host side is thin Python mirroring the reference's interfaces (EulerSolver, VocosFbank,
Vocos.decode); all compute runs in hand-written CUDA behind the C ABI declared in
``include/zipvoice_b200.h`` (built into ``zipvoice_b200/libzipvoice_b200.so``).
"""
__version__ = "0.1.0"
