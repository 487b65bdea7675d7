"""B200-native char-RNN training engine behind the reference's train/sample/data_splitter surface.

The compute path is hand-written sm_100a CUDA behind the C-ABI declared in
``include/crnn.h`` (built to ``distributed_char_rnn_b200/libcrnn.so``).  There
is no CPU fallback: importing :mod:`distributed_char_rnn_b200.model` without the
built library, or running it without a GPU, raises.
"""
__version__ = "0.1.0"
