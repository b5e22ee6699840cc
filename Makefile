# Convenience targets; the real build logic lives in akka_allreduce_b200/utils/build.py (content-hash based, in-tree).
PY ?= python

all: build

build:
	$(PY) -m akka_allreduce_b200.utils.build

rebuild:
	$(PY) -m akka_allreduce_b200.utils.build --force

test:
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu:
	$(PY) -m pytest tests -q -m gpu

tsan:
	$(PY) -m pytest tests/test_tsan.py -q

bench:
	$(PY) bench.py --gpus 1

sass:
	cuobjdump -sass akka_allreduce_b200/libaar_cuda.so | grep -oE "(LDGMC|LDG|STG|RED|ATOMG|MEMBAR)[A-Z0-9_.]*" | sort | uniq -c | sort -rn | head -30

.PHONY: all build rebuild test test-gpu tsan bench sass
