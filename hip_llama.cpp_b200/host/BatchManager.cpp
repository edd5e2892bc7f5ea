// BatchManager: slot table + KV block accounting over the batch-major KV slab
// (see include/BatchManager.hpp for how this relates to the reference's stub).
#include "BatchManager.hpp"

#include <string.h>

#include "tha_b200.h"

BatchManager::BatchManager(int batch_size_, int seq_len_, int n_layers_, int kv_dim_)
    : num_physical_blocks(0), cur_physical_blocks(0), block_size(16), seq_len(seq_len_), batch_size(batch_size_),
      n_layers(n_layers_), kv_dim(kv_dim_), nrow_blocks(0), gpu_physical_KV_blocks(nullptr),
      gpu_physical_V_blocks(nullptr), bound_(nullptr) {
  nrow_blocks = (seq_len + block_size - 1) / block_size;
  num_physical_blocks = (size_t)batch_size * nrow_blocks;
  logical_KV_table.assign(batch_size, std::make_pair(-1, 0));
  n_block_table_filled.assign(batch_size, 0);
  mask_physical_KV_block.assign(batch_size, std::vector<bool>(nrow_blocks, false));
}

BatchManager::~BatchManager() { destroy_pool(); }

// ---- paged mode ---------------------------------------------------------------------------------
bool BatchManager::create_pool(size_t n_blocks) {
  destroy_pool();
  const size_t blk_floats = (size_t)n_layers * block_size * kv_dim;
  if (n_blocks < 1) return false;
  if (tha_malloc((void**)&pool_k_, n_blocks * blk_floats * sizeof(float)) != 0) { pool_k_ = nullptr; return false; }
  if (tha_malloc((void**)&pool_v_, n_blocks * blk_floats * sizeof(float)) != 0) { tha_free(pool_k_); pool_k_ = pool_v_ = nullptr; return false; }
  if (tha_malloc((void**)&table_d_, (size_t)batch_size * nrow_blocks * sizeof(int)) != 0) {
    tha_free(pool_k_); tha_free(pool_v_); pool_k_ = pool_v_ = nullptr; table_d_ = nullptr;
    return false;
  }
  tha_memset(pool_k_, 0, n_blocks * blk_floats * sizeof(float));
  tha_memset(pool_v_, 0, n_blocks * blk_floats * sizeof(float));
  tha_memset(table_d_, 0, (size_t)batch_size * nrow_blocks * sizeof(int));
  pool_blocks_ = n_blocks;
  num_physical_blocks = n_blocks;
  cur_physical_blocks = 0;
  table_h_.assign((size_t)batch_size * nrow_blocks, 0);
  dirty_.assign(batch_size, 0);
  free_list_.clear();
  for (size_t b = n_blocks; b-- > 0;) free_list_.push_back((int)b);   // block 0 is handed out first
  for (auto& m : mask_physical_KV_block) std::fill(m.begin(), m.end(), false);
  std::fill(n_block_table_filled.begin(), n_block_table_filled.end(), 0);
  logicId_physicAddr_mapper.clear();
  gpu_physical_KV_blocks = pool_k_;
  gpu_physical_V_blocks = pool_v_;
  if (tha_kv_pool_register(pool_k_, pool_v_, table_d_, nrow_blocks, block_size, (int)n_blocks) != 0) { destroy_pool(); return false; }
  return true;
}

void BatchManager::destroy_pool() {
  if (!pool_blocks_) return;
  tha_kv_pool_unregister(pool_k_);
  if (bound_ && bound_->key_cache == pool_k_) { bound_->key_cache = nullptr; bound_->value_cache = nullptr; }
  tha_free(pool_k_); tha_free(pool_v_); tha_free(table_d_);
  pool_k_ = pool_v_ = nullptr; table_d_ = nullptr;
  pool_blocks_ = 0;
  free_list_.clear();
}

void BatchManager::bind_paged(RunState* s_d) {
  bound_ = s_d;
  if (s_d) { s_d->key_cache = pool_k_; s_d->value_cache = pool_v_; }
}

bool BatchManager::ensure_position(int seq_id, int pos) {
  if (!paged()) { return true; }
  const int need = pos / block_size;
  for (int rb = 0; rb <= need && rb < nrow_blocks; rb++) {
    if (mask_physical_KV_block[seq_id][rb]) continue;
    if (free_list_.empty()) return false;
    const int id = free_list_.back();
    free_list_.pop_back();
    mask_physical_KV_block[seq_id][rb] = true;
    n_block_table_filled[seq_id]++;
    cur_physical_blocks++;
    table_h_[(size_t)seq_id * nrow_blocks + rb] = id;
    dirty_[seq_id] = 1;
    KVBlock blk;
    blk.addr = pool_k_ + (size_t)id * n_layers * block_size * kv_dim;
    blk.occupied = 1;
    logicId_physicAddr_mapper[(size_t)id] = blk;
  }
  return true;
}

void BatchManager::sync_table(void* stream) {
  if (!paged()) return;
  for (int s = 0; s < batch_size; s++)
    if (dirty_[s]) {
      tha_memcpy_async(table_d_ + (size_t)s * nrow_blocks, table_h_.data() + (size_t)s * nrow_blocks, (size_t)nrow_blocks * sizeof(int), stream);
      dirty_[s] = 0;
    }
}

size_t BatchManager::block_bytes() const {
  return (size_t)2 * n_layers * block_size * kv_dim * sizeof(float);  // K and V rows of every layer
}

size_t BatchManager::get_gpu_memory(int d_id) {
  int cur = 0;
  tha_get_device(&cur);
  if (cur != d_id) tha_set_device(d_id);
  size_t free_b = 0, total_b = 0;
  tha_mem_info(&free_b, &total_b);
  if (cur != d_id) tha_set_device(cur);
  return free_b;
}

// How many blocks the device could hold beyond the slab already bound (admission planning).
void BatchManager::set_gpu_physical_blocks(int d_id) {
  const size_t spare = get_gpu_memory(d_id) / block_bytes();
  const size_t slab = (size_t)batch_size * nrow_blocks;
  num_physical_blocks = bound_ ? slab : (spare < slab ? spare : slab);
}

void BatchManager::bind(RunState* s_d) {
  bound_ = s_d;
  gpu_physical_KV_blocks = s_d ? s_d->key_cache : nullptr;
  gpu_physical_V_blocks = s_d ? s_d->value_cache : nullptr;
  num_physical_blocks = (size_t)batch_size * nrow_blocks;
}
void BatchManager::set_gpu_block() { bind(bound_); }

int BatchManager::acquire_slot(int request_id) {
  for (int s = 0; s < batch_size; s++)
    if (logical_KV_table[s].first < 0) {
      logical_KV_table[s] = std::make_pair(request_id, 0);
      n_block_table_filled[s] = 0;
      return s;
    }
  return -1;
}

int BatchManager::request_of(int seq_id) const { return logical_KV_table[seq_id].first; }

int BatchManager::active_slots() const {
  int n = 0;
  for (const auto& e : logical_KV_table) n += e.first >= 0;
  return n;
}

// block id = slot * nrow_blocks + row block: the slab is batch-major, so a slot's blocks are
// its own rows; "allocation" is bookkeeping that bounds the live footprint.
void BatchManager::get_KV_Block(int seq_id) {
  if (paged()) { ensure_position(seq_id, logical_KV_table[seq_id].second); return; }
  const int filled = logical_KV_table[seq_id].second;
  const int rb = filled / block_size;
  if (rb < nrow_blocks && !mask_physical_KV_block[seq_id][rb]) {
    mask_physical_KV_block[seq_id][rb] = true;
    n_block_table_filled[seq_id]++;
    cur_physical_blocks++;
    const size_t id = (size_t)seq_id * nrow_blocks + rb;
    KVBlock blk;
    blk.addr = gpu_physical_KV_blocks ? gpu_physical_KV_blocks + ((size_t)seq_id * n_layers * seq_len + (size_t)rb * block_size) * kv_dim : nullptr;
    blk.occupied = 1;
    logicId_physicAddr_mapper[id] = blk;
  }
}

void BatchManager::advance(int seq_id) {
  get_KV_Block(seq_id);
  logical_KV_table[seq_id].second++;
}

size_t BatchManager::alloc_new_block(thablasHandle_t* handle) {
  (void)handle;
  for (int s = 0; s < batch_size; s++)
    for (int rb = 0; rb < nrow_blocks; rb++)
      if (!mask_physical_KV_block[s][rb]) {
        mask_physical_KV_block[s][rb] = true;
        cur_physical_blocks++;
        const size_t id = (size_t)s * nrow_blocks + rb;
        KVBlock blk;
        blk.addr = gpu_physical_KV_blocks ? gpu_physical_KV_blocks + ((size_t)s * n_layers * seq_len + (size_t)rb * block_size) * kv_dim : nullptr;
        blk.occupied = 1;
        logicId_physicAddr_mapper[id] = blk;
        return id;
      }
  return (size_t)-1;
}

KVBlock BatchManager::get_block_by_id(thablasHandle_t* handle, size_t block_id) {
  (void)handle;
  auto it = logicId_physicAddr_mapper.find(block_id);
  if (it != logicId_physicAddr_mapper.end()) return it->second;
  KVBlock none;
  none.addr = nullptr;
  none.occupied = 0;
  return none;
}

std::vector<KVBlock> BatchManager::convert_block_ids_to_block_addrs(thablasHandle_t* handle, std::vector<size_t> block_ids) {
  std::vector<KVBlock> out;
  out.reserve(block_ids.size());
  for (size_t id : block_ids) out.push_back(get_block_by_id(handle, id));
  return out;
}

// Copy host K rows (layer 0, block_size*kv_dim floats per block, consecutive in `data`) into the blocks.
void BatchManager::push_kv_data_to_kv_cache(thablasHandle_t* handle, std::vector<size_t> kv_cache_block_ids, float* data) {
  const size_t n = (size_t)block_size * kv_dim;
  size_t off = 0;
  for (size_t id : kv_cache_block_ids) {
    KVBlock b = get_block_by_id(handle, id);
    if (b.addr) tha_memcpy_async(b.addr, data + off, n * sizeof(float), handle ? (void*)handle->copy_stream : nullptr);
    off += n;
  }
  if (handle) tha_stream_sync((void*)handle->copy_stream);
}

void BatchManager::dealloc_gpu_block(int seq_id) {
  if (paged()) {   // the slot's blocks go back to the free list (the table row keeps its stale ids: nothing reads past pos)
    for (int rb = 0; rb < nrow_blocks; rb++)
      if (mask_physical_KV_block[seq_id][rb]) {
        mask_physical_KV_block[seq_id][rb] = false;
        const int id = table_h_[(size_t)seq_id * nrow_blocks + rb];
        free_list_.push_back(id);
        logicId_physicAddr_mapper.erase((size_t)id);
        cur_physical_blocks--;
      }
    n_block_table_filled[seq_id] = 0;
    return;
  }
  for (int rb = 0; rb < nrow_blocks; rb++)
    if (mask_physical_KV_block[seq_id][rb]) {
      mask_physical_KV_block[seq_id][rb] = false;
      cur_physical_blocks--;
      logicId_physicAddr_mapper.erase((size_t)seq_id * nrow_blocks + rb);
    }
  n_block_table_filled[seq_id] = 0;
}

void BatchManager::release_slot(int seq_id) {
  dealloc_gpu_block(seq_id);
  logical_KV_table[seq_id] = std::make_pair(-1, 0);
}
