# Builds libpclip.so (the C-ABI CUDA library, sm_100a only) and the oracle's C restatement.
NVCC      ?= /usr/local/cuda/bin/nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC,-Wall,-Wno-unused-function --expt-relaxed-constexpr \
             -Xptxas -v $(EXTRA)
CSRC      := protein-clip_b200/csrc
SRCS      := $(wildcard $(CSRC)/*.cu)
OBJS      := $(patsubst $(CSRC)/%.cu,build/%.o,$(SRCS))
LIB       := protein-clip_b200/pclip/libpclip.so

all: $(LIB)

build/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh) include/pclip.h
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -cudart static -o $@ $(OBJS)

# Race / barrier / memory checks of the smoke step (needs a GPU): every kernel family once, small shapes.
# (compute-sanitizer is closed on the build pool's GPU boxes -- the target is for a maintainer's own machine.)
SANITIZER ?= /usr/local/cuda/bin/compute-sanitizer
sanitize: $(LIB)
	@for t in memcheck racecheck synccheck; do \
	  echo "== compute-sanitizer --tool $$t"; \
	  $(SANITIZER) --tool $$t --error-exitcode 1 python -c "import __graft_entry__ as g; g.smoke()" || exit 1; \
	done

clean:
	rm -rf build $(LIB)

.PHONY: all clean sanitize
