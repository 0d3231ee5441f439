// second search chain, kernels for groups of 8 lanes (see chain2_kernels.inl)
#define C2K_G8
#include "chain2_kernels.inl"
