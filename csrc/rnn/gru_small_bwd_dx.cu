// Backward of the small-hidden fused GRU family with d/dx of the network input (b200rnn_gru_bwd_dx).
#define B200_SMALL_CELL 1
#define B200_SMALL_PART 2
#define B200_SMALL_BWD_DX 1
#include "lstm_small.inc"
