// Hidden-16 members of the small-hidden fused LSTM family, forward (see lstm_small.inc; --hidden-units 16,
// src/motion/main.py:21).
#define B200_SMALL_H 16
#define B200_SMALL_PART 1
#include "lstm_small.inc"
