// plan.cuh — per-device CSR execution plan: row statistics and nnz-balanced tile tables.
//
// The reference rebuilds cuSPARSE descriptors and work buffers on every call
// (spmv.h:70-101).  Here the equivalent analysis happens once, when the matrix is placed
// on the device (CsrSparseMatrix::sync2gpu, matrix.h:331-397), and is reused by every call.
//
// A tile table partitions the merge list "rows + nnz" (Merrill & Garland's merge path,
// evaluated at row granularity): tile b owns the rows r whose merge coordinate
// f(r) = r*row_weight + rowptr[r] lies in [b*T, (b+1)*T).  f is strictly increasing, so
// the owner of each row is unique, tiles hold at most T/row_weight rows, and at most
// T + (length of their last row) nnz.  Empty-row stretches and dense stretches both cost
// ~T items per tile.
#pragma once
#include <vector>

#include "common.cuh"

namespace sblas {

struct TileTable {
  int32_t tile = 0;        // T, merge items per tile
  int32_t row_weight = 1;  // merge weight of one row
  int32_t n_tiles = 0;
  int32_t* d_tile_row = nullptr;  // [n_tiles + 1], first row of each tile; last = rows
  int32_t* d_tile_nnz = nullptr;  // [n_tiles + 1], rowptr[tile_row[b]]: first nnz of each tile; last = nnz
  int32_t max_tile_nnz = 0;
  int32_t max_tile_rows = 0;
};

}  // namespace sblas

struct sblasPlan_st {
  int device = 0;
  int32_t rows = 0;
  int64_t nnz = 0;
  const int32_t* d_rowptr = nullptr;  // not owned
  int32_t max_row_nnz = 0;
  int32_t empty_rows = 0;
  std::vector<sblas::TileTable> tables;
  // scratch for the split (carry) SpMV variant, grown on demand
  void* d_split_scratch = nullptr;
  size_t split_scratch_bytes = 0;
};

namespace sblas {

// Finds or builds the tile table (row_weight, tile) of a plan.  Builds synchronise `stream`.
int plan_get_table(sblasPlan_st* plan, cudaStream_t stream, int32_t row_weight, int32_t tile,
                   const TileTable** out);
int plan_scratch(sblasPlan_st* plan, size_t bytes, void** out);

}  // namespace sblas
