// BatchManager.hpp -- slot table + KV-block bookkeeping of the batching driver.
//
// The reference's class of this name (include/BatchManager.hpp:11-52, src/BatchManager.cpp) is
// a stub that live code never instantiates (uninitialised members, a function without a
// return, by-value mutation).  The names and members below are kept; the
// behaviour is the working version of what it gestures at:
//   * the KV cache of a GPU is one slab [batch_size][n_layers][seq_len][kv_dim] x {K, V} (the
//     layout thaDNN_s_forward_batch reads), cut into BLOCKS of `block_size` positions;
//   * a sequence slot owns the blocks that cover positions 0..pos; blocks are handed out as the
//     sequence grows and returned when the request finishes, so `cur_physical_blocks` is the
//     live KV footprint and admission control can refuse a request that would not fit;
//   * the slot table (which request sits in which slot, at which position) is the state the
//     driver loop of src/llama.cpp:968-1073 keeps in local arrays.
// PAGED mode (create_pool): the blocks are real.  K and V live in two device pools
// [n_blocks][n_layers][block_size][kv_dim]; a slot's positions are found through a row of block ids in a
// device-resident table that the attention / RoPE-scatter kernels read (tha_kv_pool_register); a block is taken
// from the free list when a sequence reaches it and returned when its request finishes, so B slots share a pool
// sized for the positions actually in flight instead of B x seq_len (Llama-2-7B fp32 KV: 2.1 GB per 2048
// positions).
#pragma once
#include <stddef.h>

#include <map>
#include <utility>
#include <vector>

#include "models.hpp"
#include "thaBLAS.hpp"

class BatchManager {
 public:
  size_t num_physical_blocks;  // capacity: batch_size * n_layers * nrow_blocks
  size_t cur_physical_blocks;  // blocks currently owned by sequences
  int block_size;              // positions per block
  int seq_len;
  int batch_size;
  int n_layers;
  int kv_dim;
  int nrow_blocks;             // blocks per (slot, layer) row = ceil(seq_len / block_size)

  std::vector<std::pair<int, int>> logical_KV_table;     // per slot: (request id or -1, positions filled)
  std::vector<int> n_block_table_filled;                 // per slot: blocks owned per layer
  std::vector<std::vector<bool>> mask_physical_KV_block; // [slot][row block] occupied
  float* gpu_physical_KV_blocks;                         // base of the K slab (device); V slab follows the same geometry
  float* gpu_physical_V_blocks;
  std::vector<float*> cpu_physical_KV_blocks;            // unused (host swap space of the reference's layer-swap variant)
  std::map<size_t, KVBlock> logicId_physicAddr_mapper;   // block id -> device address of its K rows (layer 0)

  BatchManager(int batch_size, int seq_len, int n_layers, int kv_dim);
  ~BatchManager();

  // geometry / memory
  void set_gpu_physical_blocks(int d_id);   // size the block pool from the device's free memory (informational)
  void set_gpu_block();                     // bind to a slab allocated by alloc_state_to_device_batch
  void bind(RunState* s_d);                 // same, explicit
  size_t get_gpu_memory(int d_id);          // free bytes on device d_id
  size_t block_bytes() const;               // bytes of one block (K + V, all layers)

  // slot table
  int acquire_slot(int request_id);         // first free slot or -1
  void release_slot(int seq_id);            // == dealloc_gpu_block + clear the table entry
  int request_of(int seq_id) const;
  int active_slots() const;

  // block accounting
  void get_KV_Block(int seq_id);                        // make sure the block that holds position `filled` exists
  void advance(int seq_id);                             // one more position filled (allocates at block boundaries)
  size_t alloc_new_block(thablasHandle_t* handle);      // lowest free block id, or (size_t)-1
  KVBlock get_block_by_id(thablasHandle_t* handle, size_t block_id);
  std::vector<KVBlock> convert_block_ids_to_block_addrs(thablasHandle_t* handle, std::vector<size_t> block_ids);
  void push_kv_data_to_kv_cache(thablasHandle_t* handle, std::vector<size_t> kv_cache_block_ids, float* data);
  void dealloc_gpu_block(int seq_id);

  // ---- paged mode ------------------------------------------------------------------------------
  bool create_pool(size_t n_blocks);              // allocate the pools and the device table (current device); false: no memory
  void destroy_pool();
  void bind_paged(RunState* s_d);                 // point s_d->key_cache / value_cache at the pools (the state must not own slabs)
  bool paged() const { return pool_blocks_ > 0; }
  size_t free_blocks() const { return free_list_.size(); }
  size_t blocks_for(int positions) const { return ((size_t)positions + block_size - 1) / block_size; }
  bool ensure_position(int seq_id, int pos);      // the blocks covering positions 0..pos exist (false: pool exhausted)
  void sync_table(void* stream);                  // upload table rows that changed since the last call (stream-ordered)
  const int* host_table() const { return table_h_.data(); }

 private:
  RunState* bound_;
  size_t pool_blocks_ = 0;
  float* pool_k_ = nullptr;
  float* pool_v_ = nullptr;
  int* table_d_ = nullptr;
  std::vector<int> table_h_;                      // [batch_size][nrow_blocks]
  std::vector<int> free_list_;
  std::vector<char> dirty_;                       // per slot
};
