// kernel TRI, lane vectors of 64 positions, 1 walked position above the triple (see tp_tri.cuh)
#include "tp_kernels.cuh"

#include "tp_tri.cuh"
TP_DEFINE_TRI_LAUNCHER(64, 1)
int tp_tri_gc() { return TP_TRI_GC; }
