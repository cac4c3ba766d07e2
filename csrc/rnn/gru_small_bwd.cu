// Backward half of the small-hidden fused GRU family (see gru_small_fwd.cu / lstm_small.inc).
#define B200_SMALL_CELL 1
#define B200_SMALL_PART 2
#include "lstm_small.inc"
