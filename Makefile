# Convenience targets; the Python build (python -m heat2d_b200.build) does the same.
NVCC ?= nvcc
LIBDIR := heat2d_b200/_lib
all: libs
libs:
	python -m heat2d_b200.build
example: libs
	$(NVCC) -O2 -gencode arch=compute_100a,code=sm_100a -Iinclude examples/c_api_demo.cu -L$(LIBDIR) -lheat2d_cuda -Xlinker -rpath -Xlinker '$$ORIGIN/../$(LIBDIR)' -o examples/c_api_demo
example-cpu: libs
	$(CC) -std=c99 -O2 -Iinclude examples/cpu_oracle_demo.c -L$(LIBDIR) -lheat2d_cpu -Wl,-rpath,'$$ORIGIN/../$(LIBDIR)' -o examples/cpu_oracle_demo
shim:
	python -c "import sys; sys.path.insert(0,'baseline'); import ref_build; print(ref_build.build_shim())"
test:
	python -m pytest tests -q -m "not gpu"
clean:
	rm -rf $(LIBDIR)/*.so $(LIBDIR)/obj* examples/c_api_demo examples/cpu_oracle_demo
.PHONY: all libs example example-cpu shim test clean
