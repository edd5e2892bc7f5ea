// Runtime shim: device, stream and memory calls behind the C-ABI so that the C++ host
// layer (hip_llama.cpp_b200/host) links against no CUDA library at all.
#include "common.cuh"

extern "C" int tha_device_count(int* n) { THA_CUDA_OK(cudaGetDeviceCount(n)); return THA_STATUS_SUCCESS; }
extern "C" int tha_get_device(int* dev) { THA_CUDA_OK(cudaGetDevice(dev)); return THA_STATUS_SUCCESS; }
extern "C" int tha_set_device(int dev) { THA_CUDA_OK(cudaSetDevice(dev)); return THA_STATUS_SUCCESS; }
extern "C" int tha_device_sync(void) { THA_CUDA_OK(cudaDeviceSynchronize()); return THA_STATUS_SUCCESS; }
extern "C" int tha_stream_create(void** stream) {
  cudaStream_t s;
  THA_CUDA_OK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *stream = (void*)s;
  return THA_STATUS_SUCCESS;
}
extern "C" int tha_stream_destroy(void* stream) { THA_CUDA_OK(cudaStreamDestroy((cudaStream_t)stream)); return THA_STATUS_SUCCESS; }
extern "C" int tha_stream_sync(void* stream) { THA_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream)); return THA_STATUS_SUCCESS; }
extern "C" int tha_malloc(void** dptr, size_t bytes) { THA_CUDA_OK(cudaMalloc(dptr, bytes)); return THA_STATUS_SUCCESS; }
extern "C" int tha_free(void* dptr) { THA_CUDA_OK(cudaFree(dptr)); return THA_STATUS_SUCCESS; }
extern "C" int tha_host_alloc(void** hptr, size_t bytes) { THA_CUDA_OK(cudaMallocHost(hptr, bytes)); return THA_STATUS_SUCCESS; }
extern "C" int tha_host_free(void* hptr) { THA_CUDA_OK(cudaFreeHost(hptr)); return THA_STATUS_SUCCESS; }
extern "C" int tha_memset(void* dptr, int value, size_t bytes) { THA_CUDA_OK(cudaMemset(dptr, value, bytes)); return THA_STATUS_SUCCESS; }
extern "C" int tha_memcpy_h2d(void* dst, const void* src, size_t bytes) {
  THA_CUDA_OK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice)); return THA_STATUS_SUCCESS;
}
extern "C" int tha_memcpy_d2h(void* dst, const void* src, size_t bytes) {
  THA_CUDA_OK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost)); return THA_STATUS_SUCCESS;
}
extern "C" int tha_memcpy_d2d(void* dst, const void* src, size_t bytes) {
  THA_CUDA_OK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToDevice)); return THA_STATUS_SUCCESS;
}
extern "C" int tha_memcpy_async(void* dst, const void* src, size_t bytes, void* stream) {
  THA_CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)stream)); return THA_STATUS_SUCCESS;
}
extern "C" int tha_mem_info(size_t* free_bytes, size_t* total_bytes) {
  THA_CUDA_OK(cudaMemGetInfo(free_bytes, total_bytes)); return THA_STATUS_SUCCESS;
}
