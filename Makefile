# Convenience targets; the build itself is kungfu_b200/_build.py (no CMake, no Go).
.PHONY: build test gpu-test sanitize fuzz sass docs package-check clean

build:
	python -m kungfu_b200._build

test: build
	python -m pytest tests -x -q -m "not gpu"

gpu-test: build
	python -m pytest tests -x -q -m gpu

sanitize: build
	scripts/tsan-check.sh

fuzz: build
	python scripts/fuzz_cli.py -n 200

sass: build
	python scripts/dump_sass.py

docs:
	python scripts/gen-api-doc.py

package-check:
	scripts/package-check.sh

clean:
	rm -rf build kungfu_b200/lib kungfu_b200/bin
