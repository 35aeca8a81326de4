// Device-visible context of the sharded sparse engine.  One instance per rank; the
// pointer arrays hold, for every peer p, the address of p's copy of each symmetric
// buffer as mapped into *this* process (CUDA IPC), so kernels store to / load from
// peer HBM directly over NVLink/NVSwitch.
#pragma once
#include "table.cuh"

struct PbxSparseDev {
  int rank, nranks;
  uint32_t cap;        // mailbox capacity per (src, dst) pair
  uint32_t dd_cap;     // batch dedup table capacity (power of two)
  int pull_stride;     // words per returned value row
  int push_stride;     // words per pushed gradient row
  // --- symmetric buffers (index = rank that owns the memory) ---
  uint64_t* flags[PBX_MAX_RANKS];      // [nranks] barrier epochs written by peers
  uint32_t* mail_cnt[PBX_MAX_RANKS];   // [nranks]   keys sent by src -> owner
  uint64_t* mail_keys[PBX_MAX_RANKS];  // [nranks*cap] keys sent by src -> owner
  float* pull_vals[PBX_MAX_RANKS];     // [nranks*cap*pull_stride] values owner -> requester
  float* push_mail[PBX_MAX_RANKS];     // [nranks*cap*push_stride] grads requester -> owner
  // --- local scratch ---
  uint64_t* dd_keys;    // [dd_cap]
  uint32_t* dd_uid;     // [dd_cap]
  uint32_t* send_cnt;   // [nranks] unique keys this rank routes to each owner
  uint32_t* serve_cnt;  // [nranks] keys served for each requester in the last pull
  uint32_t* mail_row;   // [nranks*cap] table row of every served key (reused by push)
  float* push_acc;      // [nranks*cap*push_stride] local merge buffer
  uint32_t* uid_multi;  // [nranks*cap] == tag of the last pull (epoch[5]) when the unique key occurred more than once in it
  uint32_t* dup_list;   // [2*max_keys] (occurrence, bag) of the non-representative occurrences of the current push
  uint8_t* occ_rep;     // [max_keys] 1 = this occurrence is the representative (first claimer) of its unique key in the last pull
  uint64_t* epoch;      // barrier epoch counter
  int* status;          // 0 ok, 1 mailbox overflow, 2 barrier timeout
  uint64_t* dbg;        // optional [grid, 16] per-CTA phase-end stamps (globaltimer), NULL = off
  int q3_tile_rows;     // push Q3: rows per warp tile (0 = generic per-thread path)
  int q3_warp_words;    // push Q3: shared-memory words per warp
  PbxTableView table;
};
