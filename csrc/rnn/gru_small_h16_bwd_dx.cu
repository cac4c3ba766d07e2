// Hidden-16 GRU backward with d/dx of the network input (input width <= 16).
#define B200_SMALL_H 16
#define B200_SMALL_CELL 1
#define B200_SMALL_PART 2
#define B200_SMALL_BWD_DX 1
#include "lstm_small.inc"
