# Convenience targets; the driver's own entry points are __graft_entry__.build() / smoke(), pytest and bench.py.
PY ?= python
.PHONY: build test test-gpu smoke bench bench-ref bindings cpp clean
build:            ## librav1e_cuda.so (nvcc, sm_100a) + oracle/libr1o.so (gcc)
	$(PY) -c "import __graft_entry__ as g; g.build()"
test: build       ## CPU suite: oracle vs golden vectors, host logic, C-ABI symbols, C / C++ / Rust boundary
	$(PY) -m pytest tests -x -q -m "not gpu"
test-gpu: build   ## parity tests proper, through the C ABI (needs a B200)
	$(PY) -m pytest tests -x -q -m gpu
smoke: build
	$(PY) -c "import __graft_entry__ as g; g.smoke()"
bench: build      ## one JSON line: 1080p 8-bit full-search ME (BASELINE.json configs[1])
	$(PY) bench.py
bench-ref: build  ## the CPU arm
	$(PY) bench.py --impl reference
bindings:         ## regenerate bindings/rust/rav1e_cuda_sys.rs from include/rav1e_cuda.h
	$(PY) tools/gen_rust_ffi.py
cpp: build        ## the reference's own unit tests through the C++ mirror (host part without a GPU; `gpu` argument on a B200)
	mkdir -p build && g++ -std=c++17 -O1 -Wall -Werror -Iinclude tests/cpp/reference_units.cpp -Lrav1e_b200 -lrav1e_cuda -Wl,-rpath,$(CURDIR)/rav1e_b200 -o build/reference_units && ./build/reference_units
clean:
	$(MAKE) -C rav1e_b200/csrc clean
	rm -rf build
