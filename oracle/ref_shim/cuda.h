// Host stand-in for <cuda.h> used only when oracle/build_ref.py compiles the reference's kernel
// string as serial C++ (test infrastructure; see oracle/__init__.py).
#pragma once
