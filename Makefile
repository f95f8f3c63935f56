# Top-level Makefile — counterpart of the reference's Makefile / aes-modes/Makefile
# (clang -O0, objects test.c + arc4.c / aes.c + aesni.c).  Here the cipher objects are
# replaced by libbulkcipher.so, whose kernels are built with
#   nvcc -gencode arch=compute_100a,code=sm_100a
# (see our-tree_b200/Makefile), and the harnesses link against it.
#
#   make            library + harnesses (our-tree_b200/build/{aes_test,rc4_test,aes_gpu_facade_test})
#   make oracle     the CPU checkers (test infrastructure only)
#   make test       CPU test suite          make gputest   GPU parity suite (needs a B200)
#   make bench      python bench.py         make clean

PY ?= python

all:
	$(MAKE) -C our-tree_b200 all

oracle:
	$(MAKE) -C oracle all

test: all oracle
	$(PY) -m pytest tests -x -q -m "not gpu"

gputest: all oracle
	$(PY) -m pytest tests -x -q -m gpu

bench: all
	$(PY) bench.py

clean:
	$(MAKE) -C our-tree_b200 clean
	$(MAKE) -C oracle clean

.PHONY: all oracle test gputest bench clean
