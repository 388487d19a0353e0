// kernel TRI, lane vectors of 80 positions, 2 walked positions above the triple (see tp_tri.cuh)
#include "tp_kernels.cuh"

#include "tp_tri.cuh"
TP_DEFINE_TRI_LAUNCHER(80, 2)
