# Builds build/libbert.so: the drop-in replacement for the reference's libbert.so (bert.h API) on top of
# hand-written sm_100a kernels.  No PyTorch, no CPU fallback.  `make oracle` builds the test-only checkers.
NVCC      ?= nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v
CXXFLAGS  := -O2 -std=c++17 -fPIC -fvisibility=hidden -Wall -Wextra -Wno-unused-parameter -I/usr/local/cuda/include
SRC       := embedding.cpp_b200/csrc
OBJ       := build/obj
CU_SRCS   := gemm_tcgen05.cu gemm_ln.cu rowwise.cu attention_tc.cu engine.cu test_entry.cu
CPP_SRCS  := api.cpp gguf.cpp tokenizer_wp.cpp quantize.cpp
OBJS      := $(addprefix $(OBJ)/,$(CU_SRCS:.cu=.o) $(CPP_SRCS:.cpp=.o))

all: build/libbert.so build/bert-server

build/libbert.so: $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -cudart static -lpthread -ldl -lrt
	ln -sf libbert.so build/libbert_shared.so

$(OBJ)/%.o: $(SRC)/%.cu $(SRC)/common.cuh $(SRC)/attention_persist.cuh $(SRC)/attention_stream.cuh $(SRC)/kernels.h include/bk_api.h | $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(OBJ)/$*.ptxas.log || (cat $(OBJ)/$*.ptxas.log; false)

$(OBJ)/%.o: $(SRC)/%.cpp $(SRC)/gguf.h $(SRC)/tokenizer_wp.h include/bert.h include/bk_api.h | $(OBJ)
	$(CXX) $(CXXFLAGS) -c $< -o $@

# the reference's examples/server.cpp protocol with dynamic batching across connections
build/bert-server: $(SRC)/server.cpp $(SRC)/server_batcher.h include/bert.h build/libbert.so
	$(CXX) $(CXXFLAGS) -o $@ $(SRC)/server.cpp -Lbuild -lbert -Wl,-rpath,'$$ORIGIN' -lpthread

$(OBJ):
	mkdir -p $(OBJ)

oracle:
	$(MAKE) -s -C oracle all

clean:
	rm -rf build

.PHONY: all oracle clean
