// Tensor-parallel context: the exchange block (flags + two partial-sum buffers) each rank owns and its peers map.
// See tp.h for the scheme; the kernels that use it are tc_prep* in step_tc.cu.
#include <string.h>

#include "common.cuh"
#include "tp.h"

using namespace tha;

struct tha_tp { TpCtx c; };

extern "C" int tha_tp_create(tha_tp_t** out, int tp, int rank, int head_size, const int* hidden_of_rank, size_t xchg_floats) {
  if (!out || tp < 1 || tp > TP_MAX || rank < 0 || rank >= tp || head_size < 2 || !hidden_of_rank || xchg_floats < 1)
    return THA_STATUS_INVALID_ARG;
  tha_tp* t = new tha_tp();
  TpCtx& c = t->c;
  c.tp = tp; c.rank = rank; c.head_size = head_size;
  for (int r = 0; r < tp; r++) {
    if (hidden_of_rank[r] < 1) { delete t; return THA_STATUS_INVALID_ARG; }
    c.hidden_of[r] = hidden_of_rank[r];
  }
  c.xchg_floats = (xchg_floats + 63) / 64 * 64;   // keeps the second buffer 256-byte aligned
  cudaError_t e = cudaGetDevice(&c.device);
  if (e == cudaSuccess) e = cudaMalloc(&c.base, c.block_bytes());
  if (e == cudaSuccess) e = cudaMemset(c.base, 0, TP_HDR_BYTES);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { set_last_error("tha_tp_create", e); delete t; return THA_STATUS_EXECUTION_FAILED; }
  c.ws.device = c.device;
  c.peer[rank] = c.base;
  c.connected = (tp == 1);
  *out = t;
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_tp_export(tha_tp_t* t, void* handle64_out) {
  if (!t || !handle64_out) return THA_STATUS_INVALID_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  THA_CUDA_OK(cudaIpcGetMemHandle(&h, t->c.base));
  memcpy(handle64_out, &h, 64);
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_tp_local_ptr(tha_tp_t* t, void** base) {
  if (!t || !base) return THA_STATUS_INVALID_ARG;
  *base = t->c.base;
  return THA_STATUS_SUCCESS;
}

extern "C" int tha_tp_connect_ipc(tha_tp_t* t, const void* handles) {
  if (!t || !handles) return THA_STATUS_INVALID_ARG;
  TpCtx& c = t->c;
  for (int r = 0; r < c.tp; r++) {
    if (r == c.rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const unsigned char*)handles + 64 * (size_t)r, 64);
    void* p = nullptr;
    THA_CUDA_OK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    c.peer[r] = (unsigned char*)p;
    c.ipc_opened[r] = true;
  }
  c.connected = true;
  return THA_STATUS_SUCCESS;
}

// Peers living in THIS process (one host thread per GPU, as the reference's driver runs: src/llama.cpp:913): plain
// device pointers; peer access is enabled here for blocks on other devices.
extern "C" int tha_tp_connect_ptr(tha_tp_t* t, void* const* bases) {
  if (!t || !bases) return THA_STATUS_INVALID_ARG;
  TpCtx& c = t->c;
  for (int r = 0; r < c.tp; r++) {
    if (!bases[r]) return THA_STATUS_INVALID_ARG;
    if (r == c.rank) continue;
    cudaPointerAttributes at;
    THA_CUDA_OK(cudaPointerGetAttributes(&at, bases[r]));
    if (at.device != c.device) {
      int can = 0;
      THA_CUDA_OK(cudaDeviceCanAccessPeer(&can, c.device, at.device));
      if (!can) return THA_STATUS_NOT_SUPPORTED;
      cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { set_last_error("cudaDeviceEnablePeerAccess", e); return THA_STATUS_EXECUTION_FAILED; }
      cudaGetLastError();
    }
    c.peer[r] = (unsigned char*)bases[r];
  }
  c.connected = true;
  return THA_STATUS_SUCCESS;
}

// mode 1 = PREPARE ONLY: the next tha_tp_decode_step_* call allocates its scratch, captures and instantiates its CUDA
// graph and caches it, but launches nothing; the next tha_tp_prefill_* call runs once without the exchange (garbage
// results; every buffer grown, every kernel loaded).  A host that drives several ranks of ONE device from one thread
// (tests/test_gpu_tp.py) prepares all ranks first: nothing that may synchronise with the device (cudaMalloc / cudaFree,
// lazy module loading, graph instantiation) then happens while a peer's kernels are already waiting for this rank.
extern "C" int tha_tp_set_mode(tha_tp_t* t, int mode) {
  if (!t || mode < 0 || mode > 1) return THA_STATUS_INVALID_ARG;
  t->c.mode = mode;
  return THA_STATUS_SUCCESS;
}

// make the rank's device current (host threads that drive several ranks)
extern "C" int tha_tp_select_device(tha_tp_t* t) {
  if (!t) return THA_STATUS_INVALID_ARG;
  THA_CUDA_OK(cudaSetDevice(t->c.device));
  return THA_STATUS_SUCCESS;
}

// 1 if a wait on a peer timed out since the last call (the step's results are then garbage, not a hang).
extern "C" int tha_tp_check_error(tha_tp_t* t) {
  if (!t) return 0;
  int v = 0;
  if (cudaMemcpy(&v, t->c.err(), sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return 1; }
  if (v) cudaMemset(t->c.err(), 0, sizeof(int));
  return v != 0;
}

extern "C" int tha_tp_destroy(tha_tp_t* t) {
  if (!t) return THA_STATUS_SUCCESS;
  TpCtx& c = t->c;
  cudaDeviceSynchronize();
  ws_release(&c.ws);
  for (int r = 0; r < c.tp; r++) if (c.ipc_opened[r]) cudaIpcCloseMemHandle(c.peer[r]);
  if (c.base) cudaFree(c.base);
  delete t;
  return THA_STATUS_SUCCESS;
}

namespace tha {
TpCtx* tp_ctx(tha_tp_t* t) { return t ? &t->c : nullptr; }
}
