// kernel SPLIT, lane vectors of 64 positions, walk depth 3 (see tp_split.cuh); one translation
// unit per instantiation so that all of them compile in parallel.
#include "tp_split.cuh"

TP_DEFINE_SPLIT_LAUNCHER(64, 3)
