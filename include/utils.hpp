// utils.hpp -- host utilities of the reference (include/utils.hpp:1-71, src/utils.cpp):
// matrix helpers used by the API tests and the checkpoint loader.
#pragma once
#include <string>

#include "models.hpp"
#include "thaBLAS.hpp"

void alloc_mat(float** m, int R, int C);
void alloc_vec(float** m, int N);
void util_free(void* m);
void rand_mat(float* m, int R, int C);  // uniform in [-0.5, 0.5)
void rand_vec(float* m, int N);
void zero_mat(float* m, int R, int C);
void zero_vec(float* m, int N);
bool compareFiles(const std::string& filePath1, const std::string& filePath2);

// v0 ("legacy") checkpoint: 7-int header, fp32 tensors (src/utils.cpp:119-178 == run.cc:110-161)
void malloc_run_state(RunState* s, Config* p);
void free_run_state(RunState* s);
void print_transformer(Transformer* t);
void memory_map_weights(TransformerWeights* w, Config* p, float* ptr, int shared_weights);
void read_checkpoint(char* checkpoint, Config* config, TransformerWeights* weights, int* fd, float** data,
                     ssize_t* file_size);
void build_transformer(Transformer* t, char* checkpoint_path);
void free_transformer(Transformer* t);

// v2 (Q8_0) checkpoint: 256-byte header, magic "ak42", int8 groups + fp32 scales (runq.c:174-251)
void build_transformer_q8(TransformerQ8* t, char* checkpoint_path);
void free_transformer_q8(TransformerQ8* t);
// 0 = legacy v0, 2 = v2 (Q8_0), -1 = unreadable
int checkpoint_version(const char* checkpoint_path);
