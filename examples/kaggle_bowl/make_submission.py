#!/usr/bin/env python
"""usage: make_submission.py sample_submission.csv test.lst extract_output.txt out.csv
Joins the text output of ``task = extract`` (one row of class probabilities per image) with the
image names of test.lst under the header of the sample submission (reference
example/kaggle_bowl/make_submission.py)."""
import csv
import sys


def main(argv):
    if len(argv) < 5:
        print("Usage: python make_submission.py sample_submission.csv test.lst text.txt out.csv")
        return 1
    with open(argv[1]) as f:
        head = next(csv.reader(f))
    with open(argv[2]) as f:
        names = [ln.rstrip("\n").split("\t")[-1].split("/")[-1] for ln in f if ln.strip()]
    with open(argv[3]) as f, open(argv[4], "w") as fo:
        w = csv.writer(fo, lineterminator="\n")
        w.writerow(head)
        for name, ln in zip(names, f):
            w.writerow([name] + ln.split())
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
