// Backward of the small-hidden fused LSTM family, variants that also emit d/dx of the network input
// (b200rnn_lstm_bwd_dx; see lstm_small.inc).  The reference never differentiates w.r.t. its input
// (src/motion/trainer/base.py:116 only steps the parameters); a drop-in nn.LSTM must.
#define B200_SMALL_PART 2
#define B200_SMALL_BWD_DX 1
#include "lstm_small.inc"
