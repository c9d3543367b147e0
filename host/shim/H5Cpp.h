// H5Cpp.h -- intentionally empty: the reference's HDF5_output.h includes "H5Cpp.h", but host/HDF5_output.h does not need
// libhdf5 (saveRankData is served by the libhdf5-free writer in libminipic_b200.so).  Present only so that sources which
// include it directly still compile.
#ifndef MINIPIC_SHIM_H5CPP_H
#define MINIPIC_SHIM_H5CPP_H
#endif
