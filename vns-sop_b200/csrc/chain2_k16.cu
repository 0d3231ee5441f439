// second search chain, kernels for groups of 16 and 32 lanes (see chain2_kernels.inl)
#include "chain2_kernels.inl"
