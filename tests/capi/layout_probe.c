/* tests/capi/layout_probe.c -- prints sizeof / offsetof of every struct and the value of every enumerator the C ABI of
 * this path shares with the reference.  Compiled twice by tests/test_cabi_reference.py:
 *   -DUSE_REFERENCE -DTI_WITH_CUDA -I/root/reference/c_api/include   -> the reference's own taichi_core.h / taichi_cuda.h
 *   -I include                                                        -> include/b200mpm.h (Part 1)
 * and the two outputs must be identical line by line. */
#include <stddef.h>
#include <stdio.h>
#ifdef USE_REFERENCE
#include <taichi/taichi.h>
#else
#include "b200mpm.h"
#endif

#define SZ(T) printf("sizeof(" #T ") = %zu\n", sizeof(T))
#define OFF(T, f) printf("offsetof(" #T ", " #f ") = %zu\n", offsetof(T, f))
#define EN(e) printf(#e " = %lld\n", (long long)(e))

int main(void) {
  EN(TI_C_API_VERSION);
  SZ(TiBool), SZ(TiFlags), SZ(TiRuntime), SZ(TiMemory), SZ(TiAotModule), SZ(TiKernel), SZ(TiComputeGraph);
  EN(TI_ERROR_SUCCESS), EN(TI_ERROR_NOT_SUPPORTED), EN(TI_ERROR_CORRUPTED_DATA), EN(TI_ERROR_NAME_NOT_FOUND), EN(TI_ERROR_INVALID_ARGUMENT);
  EN(TI_ERROR_ARGUMENT_NULL), EN(TI_ERROR_ARGUMENT_OUT_OF_RANGE), EN(TI_ERROR_ARGUMENT_NOT_FOUND), EN(TI_ERROR_INVALID_INTEROP);
  EN(TI_ERROR_INVALID_STATE), EN(TI_ERROR_INCOMPATIBLE_MODULE), EN(TI_ERROR_OUT_OF_MEMORY);
  EN(TI_ARCH_VULKAN), EN(TI_ARCH_METAL), EN(TI_ARCH_CUDA), EN(TI_ARCH_X64), EN(TI_ARCH_ARM64), EN(TI_ARCH_OPENGL), EN(TI_ARCH_GLES);
  EN(TI_DATA_TYPE_F16), EN(TI_DATA_TYPE_F32), EN(TI_DATA_TYPE_F64), EN(TI_DATA_TYPE_I8), EN(TI_DATA_TYPE_I16), EN(TI_DATA_TYPE_I32), EN(TI_DATA_TYPE_I64);
  EN(TI_DATA_TYPE_U1), EN(TI_DATA_TYPE_U8), EN(TI_DATA_TYPE_U16), EN(TI_DATA_TYPE_U32), EN(TI_DATA_TYPE_U64), EN(TI_DATA_TYPE_GEN), EN(TI_DATA_TYPE_UNKNOWN);
  EN(TI_ARGUMENT_TYPE_I32), EN(TI_ARGUMENT_TYPE_F32), EN(TI_ARGUMENT_TYPE_NDARRAY), EN(TI_ARGUMENT_TYPE_TEXTURE), EN(TI_ARGUMENT_TYPE_SCALAR), EN(TI_ARGUMENT_TYPE_TENSOR);
  EN(TI_MEMORY_USAGE_STORAGE_BIT), EN(TI_MEMORY_USAGE_UNIFORM_BIT), EN(TI_MEMORY_USAGE_VERTEX_BIT), EN(TI_MEMORY_USAGE_INDEX_BIT);
  SZ(TiMemoryAllocateInfo), OFF(TiMemoryAllocateInfo, size), OFF(TiMemoryAllocateInfo, host_write), OFF(TiMemoryAllocateInfo, host_read);
  OFF(TiMemoryAllocateInfo, export_sharing), OFF(TiMemoryAllocateInfo, usage);
  SZ(TiMemorySlice), OFF(TiMemorySlice, memory), OFF(TiMemorySlice, offset), OFF(TiMemorySlice, size);
  SZ(TiNdShape), OFF(TiNdShape, dim_count), OFF(TiNdShape, dims);
  SZ(TiNdArray), OFF(TiNdArray, memory), OFF(TiNdArray, shape), OFF(TiNdArray, elem_shape), OFF(TiNdArray, elem_type);
  SZ(TiTexture), SZ(TiScalarValue), SZ(TiScalar), OFF(TiScalar, type), OFF(TiScalar, value);
  SZ(TiTensorValue), SZ(TiTensorValueWithLength), OFF(TiTensorValueWithLength, length), OFF(TiTensorValueWithLength, data);
  SZ(TiTensor), OFF(TiTensor, type), OFF(TiTensor, contents);
  SZ(TiArgumentValue), SZ(TiArgument), OFF(TiArgument, type), OFF(TiArgument, value);
  SZ(TiNamedArgument), OFF(TiNamedArgument, name), OFF(TiNamedArgument, argument);
  SZ(TiCapabilityLevelInfo), OFF(TiCapabilityLevelInfo, capability), OFF(TiCapabilityLevelInfo, level);
  SZ(TiCudaMemoryInteropInfo), OFF(TiCudaMemoryInteropInfo, ptr), OFF(TiCudaMemoryInteropInfo, size);
  return 0;
}
