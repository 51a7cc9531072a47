# Common entry points (everything also works as the plain commands shown in README.md)
PY ?= python

.PHONY: build test test-gpu bench wheel docker tsan clean

build:            ## nvcc (sm_100a) kernel library + g++ controller core + tools, all in-tree
	$(PY) build.py

test: build       ## CPU tier: protocol, controller core, aggregation rules, stores, CKKS host, e2e over gRPC
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu: build   ## GPU tier (needs a B200)
	$(PY) -m pytest tests -q -m gpu

bench: build      ## headline round on one GPU (one JSON line)
	$(PY) bench.py

wheel:            ## runs build.py, packs the shared objects, tools, certificates and both import names
	$(PY) setup.py bdist_wheel

docker:           ## runtime image + the wheel it built in ./dist
	sh docker/build_images.sh

tsan: build       ## ThreadSanitizer build of the controller core + its protocol tests
	bash tools/sanitize.sh tsan

clean:
	rm -rf build dist *.egg-info metisfl_b200/lib/*.so metisfl_b200/_core*.so metisfl_b200/bin
