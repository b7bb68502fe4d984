# Convenience targets; the build itself is build.py (nvcc for sm_100a, g++ for the host libraries, all in-tree).
PYTHON ?= python

.PHONY: all build test test-gpu bench clean

all: build

build:
	$(PYTHON) build.py

test: build
	$(PYTHON) -m pytest tests -x -q -m "not gpu"

test-gpu: build
	$(PYTHON) -m pytest tests -x -q -m gpu

bench: build
	$(PYTHON) bench.py --gpus 1

clean:
	rm -rf build cxxnet_b200/lib/*.so bin/im2bin
