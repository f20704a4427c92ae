# Convenience targets; the driver uses __graft_entry__.build(), pytest and bench.py directly.
.PHONY: build test test-gpu bench tools clean
build:
	python -c "import __graft_entry__ as g; g.build()"
test: build
	python -m pytest tests -x -q -m "not gpu"
test-gpu: build
	python -m pytest tests -x -q -m gpu
bench: build
	python bench.py
tools: build
	$(MAKE) -C tools
clean:
	$(MAKE) -C bacchus-gpu-controller_b200/csrc clean
	$(MAKE) -C oracle clean
