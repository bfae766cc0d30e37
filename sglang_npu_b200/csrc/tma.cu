// Host-side TMA descriptor construction (driver entry point resolved at run time).
#include "sm100.cuh"

namespace b200 {

cuTensorMapEncodeTiled_t get_tensor_map_encoder() {
  static cuTensorMapEncodeTiled_t fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<cuTensorMapEncodeTiled_t>(p);
  }
  return fn;
}

int make_tmap_2d_bytes(CUtensorMap* out, const void* base, uint64_t rows, uint64_t row_bytes, uint64_t pitch_bytes,
                       uint32_t box_rows, uint32_t box_bytes, uint32_t swizzle_bytes) {
  cuTensorMapEncodeTiled_t enc = get_tensor_map_encoder();
  if (enc == nullptr) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  B200_REQUIRE(((uintptr_t)base % 16) == 0 && pitch_bytes % 16 == 0, "TMA: base and row pitch must be 16-byte aligned");
  B200_REQUIRE(box_rows >= 1 && box_rows <= 256, "TMA: box rows %u out of range", box_rows);
  cuuint64_t dims[2] = {row_bytes, rows};
  cuuint64_t strides[1] = {pitch_bytes};
  B200_REQUIRE(box_bytes >= 16 && box_bytes <= 128 && box_bytes % 16 == 0 && (swizzle_bytes == 0 || box_bytes <= swizzle_bytes),
               "TMA: inner box of %u bytes does not fit swizzle span %u", box_bytes, swizzle_bytes);
  CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  cuuint32_t box[2] = {box_bytes, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return B200_OK;
}

}  // namespace b200
