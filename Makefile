# Convenience targets; the build itself is kungfu_b200/_build.py (no CMake, no Go).
.PHONY: build test gpu-test sanitize docs clean

build:
	python -m kungfu_b200._build

test: build
	python -m pytest tests -x -q -m "not gpu"

gpu-test: build
	python -m pytest tests -x -q -m gpu

sanitize: build
	scripts/tsan-check.sh

docs:
	python scripts/gen-api-doc.py

clean:
	rm -rf build kungfu_b200/lib kungfu_b200/bin
