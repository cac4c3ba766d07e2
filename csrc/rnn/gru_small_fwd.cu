// Forward half of the small-hidden fused GRU family: the same tile machinery as the LSTM
// (lstm_small.inc) compiled with the GRU cell (GRU variants of K2/K3; the reference
// itself only builds nn.LSTM, src/motion/model.py:9).
#define B200_SMALL_CELL 1
#define B200_SMALL_PART 1
#include "lstm_small.inc"
