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
	$(PY) -m pytest tests/test_tsan.py tests/test_protocol_litmus.py -q

sanitize:
	$(PY) -m pytest tests/test_sanitizers.py -q

# default kernels' opcode sequences against the archived listings of the measured build (no GPU needed)
sass-guard:
	bench/sass_guard.sh

# CPU data plane over TCP with the contributor-count check on: a few geometries / thresholds, one line each
stress:
	bench/cpu_plane_stress.sh 28111 6 61 7 2000 0.75 2 2
	bench/cpu_plane_stress.sh 28112 8 8 1 1000 0.5 3 2
	bench/cpu_plane_stress.sh 28113 3 1000 64 2000 1.0 0 0

check: build test sass-guard stress

bench:
	$(PY) bench.py --gpus 1

sass:
	cuobjdump -sass akka_allreduce_b200/libaar_cuda.so | grep -oE "(LDGMC|LDG|STG|RED|ATOMG|MEMBAR)[A-Z0-9_.]*" | sort | uniq -c | sort -rn | head -30

.PHONY: all build rebuild test test-gpu tsan sanitize sass-guard stress check bench sass
