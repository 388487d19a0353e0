// tp_split_m96.cu -- kernel SPLIT instantiated for lane vectors of 96 positions (see tp_split.cuh).
#include "tp_split.cuh"

TP_DEFINE_SPLIT_LAUNCHER(96)
