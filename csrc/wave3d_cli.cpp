// wave3d — reference-compatible command line: wave3d N Np Lx Ly Lz [T] [timesteps]
// (ref: cuda_sol.cpp:447-475, README.txt:35-46). Everything lives in libwave3d.
#include "wave3d.h"

int main(int argc, char** argv) { return wave3d_main(argc, (const char* const*)argv); }
