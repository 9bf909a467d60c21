from decimal import ROUND_HALF_UP, Decimal

Seconds = float


def compute_num_frames(duration, frame_shift, sampling_rate):
    """Restated from memory of lhotse.utils.compute_num_frames (PARITY UNPINNED: source absent):
    the duration is snapped to a whole number of samples, then duration / frame_shift is rounded
    half up."""
    num_samples = round(duration * sampling_rate)
    window_hop = round(frame_shift * sampling_rate)
    return int(Decimal(num_samples / window_hop).quantize(0, rounding=ROUND_HALF_UP))
