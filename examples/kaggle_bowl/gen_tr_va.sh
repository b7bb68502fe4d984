#!/bin/sh
# split a shuffled list into train / validation parts (reference example/kaggle_bowl/gen_tr_va.sh)
sed -n '1,20000p' $1 > tr.lst
sed -n '20000,40000p' $1 > va.lst
