# Builds libcmbpo_b200.so (the C-ABI library) for sm_100a, in-tree.
NVCC ?= nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v
SRC := constrainedmbpo_b200/csrc
OUT := constrainedmbpo_b200/libcmbpo_b200.so
OBJS := build/ensemble_fwd.o build/wide_fwd.o build/fused_wide.o build/step_post.o build/cmbpo_abi.o build/rollout_ops.o build/gae_ops.o build/archive_ops.o build/actor_ops.o

all: $(OUT) build/umma_probe

build/%.o: $(SRC)/%.cu $(wildcard $(SRC)/*.cuh) include/cmbpo_b200.h
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(OUT): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart

build/umma_probe: tools/umma_probe.cu $(SRC)/sm100_ptx.cuh
	@mkdir -p build
	$(NVCC) $(ARCH) -O3 -lineinfo -std=c++17 -o $@ $<

# debug build with the in-kernel timeline (CMBPO_LIB=build/libcmbpo_b200_trace.so CMBPO_TRACE_OUT=file ...)
trace: build/libcmbpo_b200_trace.so
build/libcmbpo_b200_trace.so: $(wildcard $(SRC)/*.cu) $(wildcard $(SRC)/*.cuh) include/cmbpo_b200.h
	@mkdir -p build
	$(NVCC) $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -DCMBPO_TRACE -shared -o $@ $(SRC)/ensemble_fwd.cu $(SRC)/wide_fwd.cu $(SRC)/fused_wide.cu $(SRC)/step_post.cu $(SRC)/cmbpo_abi.cu $(SRC)/rollout_ops.cu $(SRC)/gae_ops.cu $(SRC)/archive_ops.cu $(SRC)/actor_ops.cu -lcudart

clean:
	rm -rf build $(OUT)
