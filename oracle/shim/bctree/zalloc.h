// oracle/shim/bctree/zalloc.h -- TEST INFRASTRUCTURE ONLY (force-included when build_const_tree.cpp is compiled, see oracle/Makefile).
// interpolate() (build_const_tree.cpp:198-330) zero-pads the coefficients only implicitly: its scratch `tmpBuff` (:202) and the
// caller's output buffer (buildConstTree :355) come straight from malloc() and rows n..nExt are never written before
// bitReverse (:287) reads them.  At production sizes those multi-gigabyte allocations are fresh zero pages from the kernel;
// at test sizes malloc() recycles heap chunks and the tool's output is garbage.  Giving the tool zeroed allocations reproduces
// its production behaviour at every size without touching its source.
#pragma once
#include <cstdlib>
#include <stdlib.h>
#define malloc(size) calloc(1, (size))
