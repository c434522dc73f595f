"""B200-native (sm_100a) implementation of the b2p2t_gru+w2v / b2p2t_gru+w2v_conformer hot path of
yuanhao-chen-nyoeghau/Wav2Vec2ForBrain.  Host code is Python/PyTorch (device memory, streams,
torch.distributed); all arithmetic on the path runs in csrc/libb2t_sm100.so through the C ABI of
include/b2t_sm100.h.  There is no CPU or library fallback."""

__all__ = ["ops", "models", "engine", "dist"]
