#!/usr/bin/env python
"""usage: gen_img_list.py train|test sample_submission.csv image_folder/ out.lst
Writes the ``index<TAB>label<TAB>path`` image list (shuffled, seed 888); same behaviour as the
reference's example/kaggle_bowl/gen_img_list.py."""
import csv
import os
import random
import sys


def main(argv):
    if len(argv) < 5:
        print("Usage: gen_img_list.py train/test sample_submission.csv train_folder img.lst")
        return 1
    random.seed(888)
    task, sample, folder, out = argv[1:5]
    with open(sample) as f:
        head = next(csv.reader(f))[1:]
    rows, cnt = [], 0
    if task == "train":
        for i, cls in enumerate(head):
            for img in sorted(os.listdir(folder + cls)):
                rows.append((cnt, i, folder + cls + "/" + img))
                cnt += 1
    else:
        for img in sorted(os.listdir(folder)):
            rows.append((cnt, 0, folder + img))
            cnt += 1
    random.shuffle(rows)
    with open(out, "w") as f:
        w = csv.writer(f, delimiter="\t", lineterminator="\n")
        for r in rows:
            w.writerow(r)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
