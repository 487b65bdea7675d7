#!/usr/bin/env python
"""Drop-in for the reference's ``sample.py`` (same flags; see distributed_char_rnn_b200/sample.py)."""
from distributed_char_rnn_b200.sample import main

if __name__ == '__main__':
    main()
