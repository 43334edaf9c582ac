# Builds the in-tree CUDA library for sm_100a (B200).  `make` here cross-compiles without a GPU.
NVCC      ?= nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v
SRCS      := $(wildcard csrc/*.cu)
OBJS      := $(patsubst csrc/%.cu,build/%.o,$(SRCS))
LIB       := fedmse_decentralized_b200/_C/libfedmse_b200.so

all: $(LIB)

build/%.o: csrc/%.cu $(wildcard csrc/*.cuh) include/fedmse_b200.h
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	@mkdir -p $(dir $(LIB))
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart

sass: $(LIB)
	@mkdir -p profiles/sass
	cuobjdump -sass $(LIB) > profiles/sass/libfedmse_b200.sass

clean:
	rm -rf build $(LIB)
.PHONY: all clean sass probes

# micro-benchmarks of the tensor-core building blocks (results quoted in profiles/README.md and DESIGN.md)
probes:
	$(NVCC) $(ARCH) -O3 -std=c++17 -Icsrc scripts/mma_probe.cu -o build/mma_probe -lcuda
	$(NVCC) $(ARCH) -O3 -std=c++17 scripts/mma_sync_probe.cu -o build/mma_sync_probe
	$(NVCC) $(ARCH) -O3 -std=c++17 -Icsrc scripts/train_step_tc_probe.cu -o build/train_step_tc_probe -lcuda
	$(NVCC) $(ARCH) -O3 -std=c++17 -Icsrc scripts/mma_rate_probe.cu -o build/mma_rate_probe
	$(NVCC) $(ARCH) -O3 -std=c++17 -Icsrc scripts/l2_to_smem_probe.cu -o build/l2_to_smem_probe
	$(NVCC) $(ARCH) -O3 -std=c++17 -Icsrc scripts/tmem_port_probe.cu -o build/tmem_port_probe
