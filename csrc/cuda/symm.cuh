// Device / host records of a symmetric (peer-mapped) memory region (see symm.cu); shared with
// the sparse push kernel, which can fold the dense-gradient allreduce into its own barriers.
#pragma once
#include "common.cuh"

struct PbxSymmDev {
  int rank, nranks;
  uint64_t* flags[PBX_MAX_RANKS];  // [64] per rank
  float* data[PBX_MAX_RANKS];      // user region per rank
  uint64_t* epoch;
  int* status;
};

struct PbxSymm {
  PbxSymmDev h;
  PbxSymmDev* d;
  bool dirty;
  void* base;
  size_t bytes, data_off;
  void* peer_base[PBX_MAX_RANKS];
  bool opened[PBX_MAX_RANKS];
  float** stat_ptrs;            // device [3]: summary vectors of the last pbx_symm_stat_update
  float* stat_ptrs_host[3];
};

