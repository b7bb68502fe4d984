#!/bin/bash
# usage: ./run.sh MNIST.conf [name=value ...]
# Trains one of the MNIST examples from this directory.  The four idx .gz files of the MNIST set belong in
# ./data/ (this machine has no network, so nothing is downloaded); when they are missing the mnist iterator
# substitutes a deterministic surrogate set of the same shape and says so.
set -e
cd "$(dirname "${BASH_SOURCE[0]}")"
mkdir -p data models
for f in train-images-idx3-ubyte.gz train-labels-idx1-ubyte.gz t10k-images-idx3-ubyte.gz t10k-labels-idx1-ubyte.gz; do
    [ -f "data/$f" ] || echo "note: data/$f not found, the surrogate set will be used"
done
exec ../../bin/cxxnet "${1:-MNIST.conf}" "${@:2}"
