"""``data_splitter.py`` of the reference (``data_splitter.py:1-36``): one global vocabulary,
then ``num_parts`` equal contiguous shards ``out_dir/data-<i>.npy`` — one per rank.

Same flags, same files, same failure (``ValueError`` from ``np.split`` when the corpus
length is not divisible by ``num_parts``).
"""
from __future__ import print_function

import argparse
import os

import numpy as np

from .utils import create_vocab_file, split_tensor


def build_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument("--data_dir", required=True, help="Data directory containing the dataset.")
    parser.add_argument("--num_parts", type=int, required=True,
                        help="Number of parts in which to divide, same as amount of worker nodes you plan to have.")
    parser.add_argument("--out_dir", default="sharded_data",
                        help="Output directory. Will contain files as  'data-<num>.npy'")
    return parser


def split(data_dir, num_parts, out_dir="sharded_data"):
    os.makedirs(out_dir, exist_ok=True)
    print("building vocabulary...")
    res = create_vocab_file(os.path.join(data_dir, "input.txt"), os.path.join(data_dir, "vocab.pkl"))
    print("sharding file...")
    shards = split_tensor(res["_ids"], num_parts)          # raises ValueError unless divisible
    paths = []
    for i, shard in enumerate(shards):
        print("writing shard %d.." % i, end="\r")
        paths.append(os.path.join(out_dir, "data-%d.npy" % i))
        np.save(paths[-1], shard)
    print("\ndone")
    return paths


def main(argv=None):
    args = build_parser().parse_args(argv)
    split(args.data_dir, args.num_parts, args.out_dir)


if __name__ == "__main__":
    main()
