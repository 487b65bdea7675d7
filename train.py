#!/usr/bin/env python
"""Drop-in for the reference's ``train.py`` (same flags; see distributed_char_rnn_b200/train.py)."""
from distributed_char_rnn_b200.train import main

if __name__ == '__main__':
    main()
