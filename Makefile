# Builds libcurvefit_cuda.so (sm_100a only) in-tree with nvcc; the CMake build
# (CMakeLists.txt) produces the same library.  `make oracle` builds the CPU
# restatement used by the tests.
NVCC     ?= nvcc
ARCH     := -gencode arch=compute_100a,code=sm_100a
NVFLAGS  := $(ARCH) -O3 -lineinfo -std=c++17 -Iinclude -Xcompiler -fPIC,-fvisibility=hidden \
            -Xptxas -v -ccbin /usr/bin/g++
# `make DEV=1`: the development switches (CF_FIT_VARIANT, CF_TREE_BITS, ... read by the sweep
# scripts under scripts/) are compiled in; a release build has no environment routes.
# The DEV library goes to scripts/dev/ (scripts/time_tree.py --dev loads it), never over the
# release one.
SRC      := $(wildcard curvefit_b200/csrc/*.cu)
ifdef DEV
NVFLAGS  += -DCF_DEVELOPMENT
OBJDIR   := build/dev
LIB      := scripts/dev/libcurvefit_cuda.so
else
OBJDIR   := build
LIB      := curvefit_b200/libcurvefit_cuda.so
endif
OBJ      := $(patsubst curvefit_b200/csrc/%.cu,$(OBJDIR)/%.o,$(SRC))

all: $(LIB)

$(OBJDIR)/%.o: curvefit_b200/csrc/%.cu curvefit_b200/csrc/common.cuh curvefit_b200/csrc/peer.cuh include/curvefit_cuda.h
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(OBJ)
	@mkdir -p $(dir $(LIB))
	$(NVCC) $(ARCH) -shared -o $@ $(OBJ) -lcudart -Xcompiler -pthread -ccbin /usr/bin/g++

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf build $(LIB)

.PHONY: all oracle clean
