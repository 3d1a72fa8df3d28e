# Top-level convenience targets (the driver uses __graft_entry__.build()).
all:
	$(MAKE) -C oni_ml_b200/csrc -j
	$(MAKE) -C oracle

test-cpu: all
	python -m pytest tests -x -q -m "not gpu"

test-gpu: all
	python -m pytest tests -x -q -m gpu

clean:
	$(MAKE) -C oni_ml_b200/csrc clean
	$(MAKE) -C oracle clean
.PHONY: all test-cpu test-gpu clean
