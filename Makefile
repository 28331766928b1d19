# Builds build/libbert.so: the drop-in replacement for the reference's libbert.so (bert.h API) on top of
# hand-written sm_100a kernels.  No PyTorch, no CPU fallback.  `make oracle` builds the test-only checkers.
NVCC      ?= nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
DEFS      ?=
NVFLAGS   := $(ARCH) $(DEFS) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v
CXXFLAGS  := -O2 -std=c++17 -fPIC -fvisibility=hidden -Wall -Wextra -Wno-unused-parameter -I/usr/local/cuda/include
SRC       := embedding.cpp_b200/csrc
BUILD     ?= build
OBJ       := $(BUILD)/obj
CU_SRCS   := gemm_tcgen05.cu gemm_ws.cu gemm_ln.cu rowwise.cu attention_tc.cu engine.cu test_entry.cu batcher_static.cu
CPP_SRCS  := api.cpp gguf.cpp tokenizer_wp.cpp quantize.cpp
OBJS      := $(addprefix $(OBJ)/,$(CU_SRCS:.cu=.o) $(CPP_SRCS:.cpp=.o))

all: $(BUILD)/libbert.so $(BUILD)/bert-server

$(BUILD)/libbert.so: $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -cudart static -lpthread -ldl -lrt
	ln -sf libbert.so $(BUILD)/libbert_shared.so

$(OBJ)/%.o: $(SRC)/%.cu $(SRC)/common.cuh $(SRC)/gemm_parts.cuh $(SRC)/attention_persist.cuh $(SRC)/attention_loop.cuh $(SRC)/attention_pack.cuh $(SRC)/attention_stream.cuh $(SRC)/attention_duo.cuh $(SRC)/attention_pt.cuh $(SRC)/attention_xs.cuh $(SRC)/kernels.h include/bk_api.h | $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(OBJ)/$*.ptxas.log || (cat $(OBJ)/$*.ptxas.log; false)

$(OBJ)/%.o: $(SRC)/%.cpp $(SRC)/gguf.h $(SRC)/tokenizer_wp.h include/bert.h include/bk_api.h | $(OBJ)
	$(CXX) $(CXXFLAGS) -c $< -o $@

# the reference's examples/server.cpp protocol with dynamic batching across connections
$(BUILD)/bert-server: $(SRC)/server.cpp $(SRC)/server_batcher.h include/bert.h $(BUILD)/libbert.so
	$(CXX) $(CXXFLAGS) -o $@ $(SRC)/server.cpp -L$(BUILD) -lbert -Wl,-rpath,'$$ORIGIN' -lpthread

$(OBJ):
	mkdir -p $(OBJ)

oracle:
	$(MAKE) -s -C oracle all

# race / memory / barrier checking of every kernel of the path (the reference has BERT_SANITIZE_* host sanitizers, CMakeLists.txt:28-30,51-64;
# here the hazards are on the device: hand-rolled mbarrier protocols, TMA / generic proxy hand-offs, TMEM aliasing).
# One tool per GPU call:  gpurun -- make sanitize TOOL=memcheck   (then racecheck, synccheck); log -> gpurun_out/sanitize_$(TOOL).log
TOOL ?= memcheck
sanitize: all
	mkdir -p gpurun_out
	timeout 1500 compute-sanitizer --tool $(TOOL) --error-exitcode 9 --launch-timeout 120 python tools/sanitize_smoke.py > gpurun_out/sanitize_$(TOOL).log 2>&1; \
	rc=$$?; tail -25 gpurun_out/sanitize_$(TOOL).log; echo "compute-sanitizer --tool $(TOOL): exit code $$rc"; exit $$rc

clean:
	rm -rf build

.PHONY: all oracle clean sanitize
