# Builds libdpt_b200.so in-tree (sm_100a only).  `make -j` ; `make sass` writes one full SASS listing per launched kernel into profiles/sass/.
NVCC      ?= /usr/local/cuda/bin/nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -Wall -Xptxas -v \
             --expt-relaxed-constexpr -Iinclude
BUILD     := build
LIB       := distributedpytorch_b200/_native/libdpt_b200.so

SRCS := $(wildcard csrc/kernels/*.cu) $(wildcard csrc/comm/*.cu) $(wildcard csrc/runtime/*.cu)
OBJS := $(patsubst csrc/%.cu,$(BUILD)/%.o,$(SRCS))
HDRS := $(wildcard csrc/common/*) include/dpt_b200.h $(wildcard csrc/runtime/*.h)

all: $(LIB)

$(BUILD)/%.o: csrc/%.cu $(HDRS)
	@mkdir -p $(dir $@)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $(BUILD)/$(subst /,_,$*).ptxas.log || (cat $(BUILD)/$(subst /,_,$*).ptxas.log; exit 1)

$(LIB): $(OBJS)
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lpthread

# microbenchmarks (scripts/micro/): cycles per tcgen05.mma vs N / operand layout / addressing
micro: $(BUILD)/mma_cost
$(BUILD)/mma_cost: scripts/micro/mma_cost.cu $(HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(ARCH) -O3 -std=c++17 -Iinclude -o $@ $<

sass: $(LIB)
	python scripts/dump_sass.py

clean:
	rm -rf $(BUILD) $(LIB)

.PHONY: all clean sass micro
