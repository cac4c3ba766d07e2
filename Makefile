# Convenience targets; everything here is a one-line wrapper around a command documented in README.md / INTEGRATION.md.
PY ?= python
NPROC ?= auto

.PHONY: build rebuild test test-gpu test-multigpu smoke bench bench-reference bench-scaled train wheel sass clean

build:            ## compile csrc/ for sm_100a into pytorch_distributed_rnn_b200/_lib/libb200rnn.so (only if stale)
	$(PY) -m pytorch_distributed_rnn_b200._build

rebuild:
	$(PY) -m pytorch_distributed_rnn_b200._build --force

test: build       ## CPU suite (gloo, world size 2)
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build   ## on a B200
	$(PY) -m pytest tests -x -q -m gpu

test-multigpu: build   ## on >= 2 B200s
	$(PY) -m pytest tests -x -q -m multigpu

smoke: build
	$(PY) -c "import __graft_entry__ as g; g.smoke()"

bench: build      ## headline metric, one GPU; N GPUs: torchrun --nproc-per-node N bench.py --gpus N
	$(PY) bench.py

bench-reference:  ## the unmodified reference on the same config (installs it under baseline/_ref on first use)
	$(PY) bench.py --impl reference

bench-scaled: build
	$(PY) bench.py --config scaled

train:            ## one epoch of the motion model on all visible GPUs, synthetic data
	$(PY) -m pytorch_distributed_rnn_b200.parallel.launch -n $(NPROC) -m pytorch_distributed_rnn_b200 --synthetic --epochs 1 distributed

wheel: build
	sh deploy/build_wheel.sh

sass: build       ## opcode histogram + listings under profiles/sass/ (no GPU needed)
	$(PY) tools/sass_histogram.py

clean:
	rm -rf build dist pytorch_distributed_rnn_b200/_lib/*.so
