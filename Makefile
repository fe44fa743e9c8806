# Builds mr_vamp_b200/lib/libvamp_b200.so (hand-written sm_100a kernels + the C ABI of include/vamp_b200.h).
#   make            generate robot traces from $(VAMP_SRC) when it is present, compile, link
#   make oracle     test infrastructure only: the reference compiled behind oracle/ref/vref_api.h
VAMP_SRC ?= /root/reference
PYTHON ?= python
NVCC ?= nvcc
PKG := mr_vamp_b200
GEN := $(PKG)/_gen
OBJ := build/obj
LIB := $(PKG)/lib/libvamp_b200.so
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := -std=c++17 -O3 -diag-suppress 177 $(ARCH) -lineinfo -Xcompiler -fPIC -Xptxas -v
ROBOTS := sphere ur5 panda fetch baxter
CSRC := $(PKG)/csrc
HDRS := $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.h) include/vamp_b200.h

cap = $(shell echo $(1) | sed 's/./\U&/')

PYEXT := $(PKG)/_core_ext$(shell $(PYTHON) -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
PYINC := $(shell $(PYTHON) -c "import pybind11, sysconfig; print('-I' + pybind11.get_include(), '-I' + sysconfig.get_paths()['include'])")

all: $(LIB) $(PYEXT)

# compiled Python module on the surface of the reference's nanobind module (pybind11; host code only, links the C ABI)
$(PYEXT): $(CSRC)/vb_pymodule.cc include/vamp_b200.h $(LIB)
	g++ -std=c++17 -O2 -fPIC -shared -fvisibility=hidden $(PYINC) $< -L$(PKG)/lib -lvamp_b200 -Wl,-rpath,'$$ORIGIN/lib' -o $@

# The traces are generated from the VAMP checkout the library is built against; without a checkout
# (e.g. on a GPU box that received prebuilt files) existing generated headers are used as they are.
$(GEN)/.stamp: $(wildcard $(PKG)/codegen/*.py)
	@if [ -d $(VAMP_SRC)/src/impl/vamp ]; then VAMP_SRC=$(VAMP_SRC) $(PYTHON) -m $(PKG).codegen.emit_cuda >/dev/null; \
	 elif [ ! -f $(GEN)/panda_trace.cuh ]; then echo "no VAMP checkout at $(VAMP_SRC) and no generated traces"; exit 1; fi
	@mkdir -p $(GEN) && touch $@

$(OBJ)/vb_robot_%.o: $(CSRC)/vb_robot.cu $(GEN)/.stamp $(HDRS)
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -DVB_TRACE_HEADER='"../_gen/$*_trace.cuh"' -DVB_TRACE=$(call cap,$*)Trace -DVB_OPS=robot_ops_$* \
	    -c $< -o $@ 2> $(OBJ)/vb_robot_$*.ptxas.log || (cat $(OBJ)/vb_robot_$*.ptxas.log; exit 1)

$(OBJ)/vb_capi.o: $(CSRC)/vb_capi.cu $(HDRS)
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(OBJ)/vb_capi.ptxas.log || (cat $(OBJ)/vb_capi.ptxas.log; exit 1)

$(OBJ)/vb_capt_build.o: $(CSRC)/vb_capt_build.cu $(CSRC)/vb_capt_build.h
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(OBJ)/vb_capt_build.ptxas.log || (cat $(OBJ)/vb_capt_build.ptxas.log; exit 1)

$(LIB): $(OBJ)/vb_capi.o $(OBJ)/vb_capt_build.o $(foreach r,$(ROBOTS),$(OBJ)/vb_robot_$(r).o)
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -shared -o $@ $^ -ldl

oracle:
	$(MAKE) -C oracle VAMP_SRC=$(VAMP_SRC) _ref/libvamp_ref.so

clean:
	rm -rf build $(PKG)/lib $(GEN)

.PHONY: all oracle clean
