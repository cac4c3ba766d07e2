// Hidden-16 members of the small-hidden fused LSTM family, backward.
#define B200_SMALL_H 16
#define B200_SMALL_PART 2
#include "lstm_small.inc"
