// Hidden-16 members of the small-hidden fused GRU family, forward.
#define B200_SMALL_H 16
#define B200_SMALL_CELL 1
#define B200_SMALL_PART 1
#include "lstm_small.inc"
