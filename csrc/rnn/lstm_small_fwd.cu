// Forward half of the small-hidden fused LSTM family (see lstm_small.inc); split into two translation
// units so the ~50 kernel instantiations compile in parallel.
#define B200_SMALL_PART 1
#include "lstm_small.inc"
