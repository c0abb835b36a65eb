// Host-side TMA tensor-map construction (driver API cuTensorMapEncodeTiled).
#include "common.cuh"
#include "tc_common.cuh"

namespace opb {

typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                              const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// The driver entry point is resolved at run time so that the library loads (and exports its ABI) on hosts
// without libcuda.so.1 -- nothing here links against the driver library.
static encode_fn get_encoder() {
    static encode_fn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
            set_error("cuTensorMapEncodeTiled is not available from the CUDA driver (%s)", cudaGetErrorString(e));
            return nullptr;
        }
        encode = (encode_fn)fn;
    }
    return encode;
}

static int encode_3d(CUtensorMap *map, CUtensorMapDataType dt, int elem_bytes, const void *base, int batch, int rows, int cols,
                     int box_cols, int box_rows, CUtensorMapSwizzle swz) {
    encode_fn encode = get_encoder();
    if (!encode) return OPB_ERR_CUDA;
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)batch};
    cuuint64_t strides[2] = {(cuuint64_t)cols * elem_bytes, (cuuint64_t)rows * cols * elem_bytes};     // bytes, dims 1 and 2
    cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(map, dt, 3, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed: CUresult %d (rows=%d cols=%d batch=%d elem=%d)", (int)r, rows, cols, batch, elem_bytes);
        return OPB_ERR_CUDA;
    }
    return OPB_OK;
}

int make_tmap_f16_3d(CUtensorMap *map, const void *base, int batch, int rows, int cols, int box_rows) {
    return encode_3d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, batch, rows, cols, 64, box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_f16_3d_k32(CUtensorMap *map, const void *base, int batch, int rows, int cols, int box_rows) {
    return encode_3d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, batch, rows, cols, 32, box_rows, CU_TENSOR_MAP_SWIZZLE_64B);
}

int make_tmap_u8_3d(CUtensorMap *map, const void *base, int batch, int rows, int cols, int box_rows) {
    return encode_3d(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, base, batch, rows, cols, 128, box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_f32_3d_store(CUtensorMap *map, const void *base, int batch, int rows, int cols, int box_rows) {
    return encode_3d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, batch, rows, cols, 16, box_rows, CU_TENSOR_MAP_SWIZZLE_64B);
}

int make_tmap_f32_3d_store_wide(CUtensorMap *map, const void *base, int batch, int rows, int cols, int box_rows) {
    return encode_3d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, batch, rows, cols, 32, box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_u8_3d_store(CUtensorMap *map, const void *base, int batch, int rows, int cols, int box_rows) {
    return encode_3d(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, base, batch, rows, cols, 64, box_rows, CU_TENSOR_MAP_SWIZZLE_64B);
}

int make_tmap_f16_3d_store(CUtensorMap *map, const void *base, int batch, int rows, int cols, int box_rows) {
    return encode_3d(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, batch, rows, cols, 32, box_rows, CU_TENSOR_MAP_SWIZZLE_64B);
}

static int encode_nhwc(CUtensorMap *map, const void *base, int B, int H, int W, int C, int box_c, int box_w, int box_h, CUtensorMapSwizzle swz) {
    encode_fn encode = get_encoder();
    if (!encode) return OPB_ERR_CUDA;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {(cuuint32_t)box_c, (cuuint32_t)box_w, (cuuint32_t)box_h, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void *>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (NHWC) failed: CUresult %d (B=%d H=%d W=%d C=%d box %dx%d)", (int)r, B, H, W, C, box_w, box_h);
        return OPB_ERR_CUDA;
    }
    return OPB_OK;
}

int make_tmap_f16_nhwc(CUtensorMap *map, const void *base, int B, int H, int W, int C, int box_w, int box_h) {
    return encode_nhwc(map, base, B, H, W, C, 64, box_w, box_h, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_f16_nhwc_store(CUtensorMap *map, const void *base, int B, int H, int W, int C, int box_w, int box_h) {
    return encode_nhwc(map, base, B, H, W, C, 32, box_w, box_h, CU_TENSOR_MAP_SWIZZLE_64B);
}

}  // namespace opb
