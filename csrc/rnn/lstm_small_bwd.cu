// Backward half of the small-hidden fused LSTM family (see lstm_small.inc).
#define B200_SMALL_PART 2
#include "lstm_small.inc"
