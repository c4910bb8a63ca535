// TEST INFRASTRUCTURE (oracle). Not part of the product path.
//
// cuda_shim.h -- the minimum CUDA surface needed to compile the reference's
// DEVICE sources (/root/reference/src/raytracer.cu and everything it
// #includes: objects.cu, ray.cu, camera.cu, material.cu, matrix.cu, utils.cu)
// as ordinary host C++ with g++.  Nothing here restates reference arithmetic:
// the reference source itself is what gets compiled.
//
// Semantics chosen to match what nvcc gives the same source on the device:
//  * sqrt/log/cos/sin/asin/acos/tan on a float argument resolve to the float
//    functions (CUDA overloads) -- <math.h> in C++ mode provides those;
//  * min/max are the CUDA global overloads, including the mixed
//    double min(float, double) that keeps ray.cu:101-102 in FP64;
//  * pow(float, int) returns float (CUDA's powif-style overload).
#pragma once
#include <string>
#include <vector>
#include <stdexcept>
#include <fstream>
#include <chrono>
#include <random>
#include <utility>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include <math.h>
#include <cmath>

#define __host__
#define __device__
#define __global__
#define __constant__
#define __forceinline__ inline

typedef unsigned int uint;

struct float3 { float x, y, z; };

struct dim3 {
    unsigned int x, y, z;
    dim3(unsigned int a = 1, unsigned int b = 1, unsigned int c = 1) : x(a), y(b), z(c) {}
};
struct uint3_shim { unsigned int x, y, z; };
static thread_local uint3_shim threadIdx = {0, 0, 0};
static thread_local uint3_shim blockIdx = {0, 0, 0};
static thread_local dim3 blockDim(1, 1, 1);
static thread_local dim3 gridDim(1, 1, 1);

typedef int cudaError_t;
static const cudaError_t cudaSuccess = 0;
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };

static inline const char *cudaGetErrorString(cudaError_t) { return "host shim"; }
static inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaMalloc(void **p, size_t n) { *p = malloc(n ? n : 1); return *p ? cudaSuccess : 2; }
template <typename T> static inline cudaError_t cudaMallocManaged(T **p, size_t n) { *p = (T *)malloc(n ? n : 1); return *p ? cudaSuccess : 2; }
static inline cudaError_t cudaFree(void *p) { free(p); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { if (n) memcpy(d, s, n); return cudaSuccess; }
template <typename T> static inline cudaError_t cudaMemcpyToSymbol(T &sym, const void *s, size_t n) { memcpy((void *)&sym, s, n); return cudaSuccess; }

// CUDA's global min/max overload set (the ones the reference resolves to).
static inline float min(float a, float b) { return fminf(a, b); }
static inline float max(float a, float b) { return fmaxf(a, b); }
static inline double min(double a, double b) { return fmin(a, b); }
static inline double max(double a, double b) { return fmax(a, b); }
static inline double min(float a, double b) { return fmin((double)a, b); }
static inline double min(double a, float b) { return fmin(a, (double)b); }
static inline double max(float a, double b) { return fmax((double)a, b); }
static inline double max(double a, float b) { return fmax(a, (double)b); }
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
// float max(float, int): objects.cu:407 uses tmax = INF (int) then min(tmax, float)
static inline float min(float a, int b) { return fminf(a, (float)b); }
static inline float max(float a, int b) { return fmaxf(a, (float)b); }
static inline float min(int a, float b) { return fminf((float)a, b); }
static inline float max(int a, float b) { return fmaxf((float)a, b); }

// CUDA: float pow(float, int)
static inline float pow(float a, int b) { return powf(a, (float)b); }
