"""``sample.py`` of the reference (``sample.py:13-46``): same flags, same files, same output."""
from __future__ import print_function

import argparse
import os
import pickle

from . import checkpoint as ckpt_io


def build_parser():
    p = argparse.ArgumentParser(formatter_class=argparse.ArgumentDefaultsHelpFormatter)
    p.add_argument('--save_dir', type=str, default='save', help='model directory to store checkpointed models')
    p.add_argument('-n', type=int, default=500, help='number of characters to sample')
    p.add_argument('--prime', type=str, default=u'', help='prime text')
    p.add_argument('--sample', type=int, default=1,
                   help='0 to use max at each timestep, 1 to sample at each timestep, 2 to sample on spaces')
    return p


def sample(args):
    with open(os.path.join(args.save_dir, 'config.pkl'), 'rb') as f:
        saved_args = pickle.load(f)
    with open(os.path.join(args.save_dir, 'chars_vocab.pkl'), 'rb') as f:
        chars, vocab = pickle.load(f)
    if args.prime == '':                    # sample.py:36-37: most frequent char
        args.prime = chars[0]
    from .model import Model
    saved_args.distributed = False
    model = Model(saved_args, training=False, precision=getattr(saved_args, "precision", "auto"))
    ckpt = ckpt_io.latest_checkpoint(args.save_dir)
    if ckpt:                                # sample.py:42-46: silent when there is no checkpoint
        model.load_state_dict(ckpt_io.load(ckpt))
        text = model.sample(None, chars, vocab, args.n, args.prime, args.sample)
        print(text.encode('utf-8'))
        return text
    return None


def main(argv=None):
    sample(build_parser().parse_args(argv))


if __name__ == '__main__':
    main()
