// Register-level combine of 16-byte vectors for every dtype x ReduceOp the API accepts
// (main.py:14-15,23-24 name SUM/PRODUCT/MAX/MIN). 16-bit floats are widened to fp32 for the
// whole combine and rounded once, so a P2P reduction over N peers is as accurate as NVLS
// with acc::f32. Peers are always combined in rank order 0..N-1, which is what makes the
// result bitwise identical on every rank.
#pragma once

#include <type_traits>

#include "device.cuh"

namespace pdcc {

template <typename T>
struct Traits;

#define PDCC_TRAITS(TYPE, COMP, ENUMV)                                   \
  template <>                                                            \
  struct Traits<TYPE> {                                                  \
    using C = COMP;                                                      \
    static constexpr int kEnum = ENUMV;                                  \
    static constexpr int K = 16 / sizeof(TYPE);                          \
    __device__ static __forceinline__ C up(TYPE v) { return static_cast<C>(v); } \
    __device__ static __forceinline__ TYPE down(C v) { return static_cast<TYPE>(v); } \
  };
PDCC_TRAITS(float, float, PDCC_F32)
PDCC_TRAITS(double, double, PDCC_F64)
PDCC_TRAITS(int32_t, int32_t, PDCC_I32)
PDCC_TRAITS(int64_t, int64_t, PDCC_I64)
PDCC_TRAITS(uint8_t, int32_t, PDCC_U8)
PDCC_TRAITS(int8_t, int32_t, PDCC_I8)
PDCC_TRAITS(int16_t, int32_t, PDCC_I16)
#undef PDCC_TRAITS

template <>
struct Traits<__nv_bfloat16> {
  using C = float;
  static constexpr int kEnum = PDCC_BF16;
  static constexpr int K = 8;
  __device__ static __forceinline__ float up(__nv_bfloat16 v) { return __bfloat162float(v); }
  __device__ static __forceinline__ __nv_bfloat16 down(float v) { return __float2bfloat16_rn(v); }
};
template <>
struct Traits<__half> {
  using C = float;
  static constexpr int kEnum = PDCC_F16;
  static constexpr int K = 8;
  __device__ static __forceinline__ float up(__half v) { return __half2float(v); }
  __device__ static __forceinline__ __half down(float v) { return __float2half_rn(v); }
};

template <int OP, typename C>
__device__ __forceinline__ C combine(C a, C b) {
  if constexpr (OP == PDCC_SUM) return a + b;
  else if constexpr (OP == PDCC_PRODUCT) return a * b;
  else if constexpr (OP == PDCC_MIN) {
    if constexpr (std::is_same<C, float>::value) return fminf(a, b);
    else if constexpr (std::is_same<C, double>::value) return fmin(a, b);
    else return a < b ? a : b;
  } else {
    if constexpr (std::is_same<C, float>::value) return fmaxf(a, b);
    else if constexpr (std::is_same<C, double>::value) return fmax(a, b);
    else return a > b ? a : b;
  }
}

template <typename C>
__device__ __forceinline__ C apply_scale(C v, float scale) {
  if constexpr (std::is_integral<C>::value) {
    return static_cast<C>(static_cast<double>(v) * static_cast<double>(scale));
  } else {
    return v * static_cast<C>(scale);
  }
}

// Accumulator for one 16-byte unit of T.
template <typename T>
struct Acc {
  using Tr = Traits<T>;
  using C = typename Tr::C;
  C v[Tr::K];

  __device__ __forceinline__ void init(const uint4& u) {
    const T* e = reinterpret_cast<const T*>(&u);
#pragma unroll
    for (int i = 0; i < Tr::K; ++i) v[i] = Tr::up(e[i]);
  }
  template <int OP>
  __device__ __forceinline__ void add(const uint4& u) {
    const T* e = reinterpret_cast<const T*>(&u);
#pragma unroll
    for (int i = 0; i < Tr::K; ++i) v[i] = combine<OP, C>(v[i], Tr::up(e[i]));
  }
  __device__ __forceinline__ void scale(float s) {
#pragma unroll
    for (int i = 0; i < Tr::K; ++i) v[i] = apply_scale<C>(v[i], s);
  }
  __device__ __forceinline__ uint4 pack() const {
    uint4 u;
    T* e = reinterpret_cast<T*>(&u);
#pragma unroll
    for (int i = 0; i < Tr::K; ++i) e[i] = Tr::down(v[i]);
    return u;
  }
};

// Write the K results of one unit to the user tensor. `valid` elements of the unit exist
// (the last unit of a message may be ragged). WIDE: T is a 16-bit float and the output
// tensor is fp32 (the DP demo's fused cast, README.md:5 / BASELINE.json config 5).
template <typename T, bool WIDE>
__device__ __forceinline__ void store_unit(void* out_base, size_t elem_index, const Acc<T>& acc,
                                           int valid) {
  using Tr = Traits<T>;
  if constexpr (WIDE) {
    float* o = static_cast<float*>(out_base) + elem_index;
    if (valid == Tr::K && aligned16(o)) {
#pragma unroll
      for (int i = 0; i < Tr::K; i += 4) {
        uint4 u;
        float* f = reinterpret_cast<float*>(&u);
        f[0] = float(acc.v[i]); f[1] = float(acc.v[i + 1]);
        f[2] = float(acc.v[i + 2]); f[3] = float(acc.v[i + 3]);
        st_v4(o + i, u);
      }
    } else {
      for (int i = 0; i < valid; ++i) o[i] = float(acc.v[i]);
    }
  } else {
    T* o = static_cast<T*>(out_base) + elem_index;
    if (valid == Tr::K && aligned16(o)) {
      st_v4(o, acc.pack());
    } else {
      uint4 u = acc.pack();
      const T* e = reinterpret_cast<const T*>(&u);
      for (int i = 0; i < valid; ++i) o[i] = e[i];
    }
  }
}

// Which multimem kind (if any) implements (T, OP) inside the switch.
template <typename T, int OP>
struct MmFor {
  static constexpr int kind = MM_NONE;
};
#define PDCC_MMFOR(TYPE, OPV, KINDV) \
  template <>                        \
  struct MmFor<TYPE, OPV> {          \
    static constexpr int kind = KINDV; \
  };
PDCC_MMFOR(float, PDCC_SUM, MM_ADD_F32)
PDCC_MMFOR(double, PDCC_SUM, MM_ADD_F64)
PDCC_MMFOR(__nv_bfloat16, PDCC_SUM, MM_ADD_BF16)
PDCC_MMFOR(__nv_bfloat16, PDCC_MIN, MM_MIN_BF16)
PDCC_MMFOR(__nv_bfloat16, PDCC_MAX, MM_MAX_BF16)
PDCC_MMFOR(__half, PDCC_SUM, MM_ADD_F16)
PDCC_MMFOR(__half, PDCC_MIN, MM_MIN_F16)
PDCC_MMFOR(__half, PDCC_MAX, MM_MAX_F16)
PDCC_MMFOR(int32_t, PDCC_SUM, MM_ADD_S32)
PDCC_MMFOR(int32_t, PDCC_MIN, MM_MIN_S32)
PDCC_MMFOR(int32_t, PDCC_MAX, MM_MAX_S32)
PDCC_MMFOR(int64_t, PDCC_SUM, MM_ADD_U64)
PDCC_MMFOR(int64_t, PDCC_MIN, MM_MIN_S64)
PDCC_MMFOR(int64_t, PDCC_MAX, MM_MAX_S64)
#undef PDCC_MMFOR

}  // namespace pdcc
