"""B200-native Stage-1 projector-alignment hot path (drop-in for
SabaPivot/ProjectionTrainer Stage1).  Compute lives in libpt_stage1.so
(hand-written sm_100a CUDA behind the C ABI in include/pt_stage1.h)."""

__version__ = "0.1.0"
