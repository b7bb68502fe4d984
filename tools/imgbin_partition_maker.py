#!/usr/bin/env python
"""Write a Makefile that shards an image list into <= N MB ``.lst`` / ``.bin`` partitions packed by
``im2bin`` (same options and output as the reference's ``tools/imgbin-partition-maker.py``)."""
import argparse
import os
import random
import sys


def main(argv=None):
    random.seed(888)
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
        lst = [ln for ln in f if ln.strip()]
    if args.shuffle == "1":
        random.shuffle(lst)
    out = args.out if args.out.endswith("/") else args.out + "/"
    os.makedirs(out, exist_ok=True)
    limit = int(args.partition_size) << 20
    objs, cmds = [], []
    fw, sz, img_cnt, cnt = None, 0, 1, 1
    for item in lst:
        if fw is None or sz + 10240 > limit:
            lst_name = out + (args.prefix % cnt) + ".lst"
            bin_name = out + (args.prefix % cnt) + ".bin"
            objs.append(bin_name)
            if fw is not None:
                fw.close()
            fw = open(lst_name, "w")
            cmds.append("%s: %s\n\t%s %s %s %s" % (bin_name, lst_name, args.im2bin, lst_name, args.img_root, bin_name))
            sz, img_cnt = 0, 1
            cnt += 1
        path = item.rstrip("\n").split("\t")[2]
        sz += os.path.getsize(args.img_root + path) + (img_cnt + 2) * 4
        fw.write(item)
        img_cnt += 1
    if fw is not None:
        fw.close()
    with open(args.makefile, "w") as fo:
        fo.write("all: " + " ".join(objs) + "\n")
        fo.write("\n\n".join(cmds) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
