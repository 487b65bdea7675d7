#!/usr/bin/env python
"""Drop-in for the reference's ``data_splitter.py`` (same flags, same outputs)."""
from distributed_char_rnn_b200.data_splitter import main

if __name__ == "__main__":
    main()
