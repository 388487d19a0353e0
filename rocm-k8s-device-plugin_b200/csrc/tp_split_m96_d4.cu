// kernel SPLIT, lane vectors of 96 positions, walk depth 4 (see tp_split.cuh); one translation
// unit per instantiation so that all of them compile in parallel.
#include "tp_split.cuh"

TP_DEFINE_SPLIT_LAUNCHER(96, 4)
