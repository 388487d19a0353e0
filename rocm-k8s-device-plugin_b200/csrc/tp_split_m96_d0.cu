// kernel SPLIT, lane vectors of 96 positions, walk depth 0 (see tp_split.cuh); one translation
// unit per instantiation so that the 15 of them compile in parallel.
#include "tp_split.cuh"

TP_DEFINE_SPLIT_LAUNCHER(96, 0)
