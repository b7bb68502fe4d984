#!/usr/bin/env python
"""Shard an image list into partitions of bounded packed size and write the Makefile that packs every
partition with ``im2bin`` (run it with ``make -f Gen.mk -j``).  Same command line as the reference's
``tools/imgbin-partition-maker.py``; the ``imgbin`` iterator reads the parts back through
``image_conf_prefix`` / ``image_conf_ids``.

A partition is closed when the next image would push its packed size (image bytes plus the 4-byte entries of
the page's offset table) past ``--partition_size`` MB, so no ``.bin`` exceeds the bound."""
import argparse
import os
import random
import sys


def plan_partitions(entries, sizes, limit_bytes):
    """greedy split of `entries` (list lines) into consecutive groups whose packed size stays <= limit."""
    groups, cur, used = [], [], 0
    for line, nbytes in zip(entries, sizes):
        cost = nbytes + 4                      # blob + its slot in the offset table
        if cur and used + cost + 8 > limit_bytes:
            groups.append(cur)
            cur, used = [], 0
        cur.append(line)
        used += cost
    if cur:
        groups.append(cur)
    return groups


def main(argv=None):
    ap = argparse.ArgumentParser(description="Generate a Makefile to make partition imgbin files")
    ap.add_argument("--img_list", required=True, help="path to list of all images")
    ap.add_argument("--img_root", required=True, help="prefix path to the file path in img_list")
    ap.add_argument("--im2bin", default="./im2bin", help="path to im2bin tool")
    ap.add_argument("--partition_size", default="256", help="max size of a single bin file (MB)")
    ap.add_argument("--shuffle", default="0", help="shuffle the list or not")
    ap.add_argument("--prefix", required=True, help="printf-style prefix of output lists and bins, e.g. part%%03d")
    ap.add_argument("--out", required=True, help="output folder for image bins and lists")
    ap.add_argument("--makefile", default="Gen.mk", help="name of the generated Makefile")
    args = ap.parse_args(argv)

    with open(args.img_list) as f:
        entries = [ln if ln.endswith("\n") else ln + "\n" for ln in f if ln.strip()]
    if args.shuffle == "1":
        random.Random(888).shuffle(entries)
    sizes = [os.path.getsize(args.img_root + e.rstrip("\n").split("\t")[2]) for e in entries]
    out_dir = os.path.join(args.out, "")
    os.makedirs(out_dir, exist_ok=True)

    rules = []
    for number, group in enumerate(plan_partitions(entries, sizes, int(args.partition_size) << 20), start=1):
        stem = out_dir + (args.prefix % number)
        with open(stem + ".lst", "w") as fl:
            fl.writelines(group)
        rules.append((stem + ".bin", stem + ".lst"))
    with open(args.makefile, "w") as mk:
        mk.write("all: %s\n" % " ".join(target for target, _ in rules))
        for target, source in rules:
            mk.write("\n%s: %s\n\t%s %s %s %s\n" % (target, source, args.im2bin, source, args.img_root, target))
    return 0


if __name__ == "__main__":
    sys.exit(main())
