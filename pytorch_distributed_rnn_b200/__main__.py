"""``python -m pytorch_distributed_rnn_b200 ...`` = the reference's ``python main.py ...`` (src/motion/main.py:79-80)."""
from .cli import main

if __name__ == "__main__":
    main()
