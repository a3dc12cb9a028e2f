# Convenience targets (the driver uses __graft_entry__.build(), pytest and bench.py directly).
.PHONY: build test test-gpu bench clean
build:
	$(MAKE) -C quantumcollocation.jl_b200/csrc
	$(MAKE) -C oracle
test: build
	python -m pytest tests -x -q -m "not gpu"
test-gpu: build
	python -m pytest tests -x -q -m gpu
bench: build
	python bench.py --gpus 1 --steps 5 --warmup 3
clean:
	$(MAKE) -C quantumcollocation.jl_b200/csrc clean
	$(MAKE) -C oracle clean
