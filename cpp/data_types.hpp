// cpp/data_types.hpp — element type, stream type and communicator type of the b200 build.
//
// Counterpart of the reference's cpp/data_types.hpp:29-131. Differences, on purpose:
//  * the reference's element type is muddled (`_FLOAT` is float in every shipped build while
//    NCCL is told bf16, data_types.hpp:37,49-51): here bytes allocated == bytes on the wire.
//    `_FLOAT` is the compile-time DEFAULT (-DPROXY_FLOAT32 / -DPROXY_BFLOAT16, else fp16);
//    the drivers also take `-t fp32|fp16|bf16` at run time.
//  * host code never includes CUDA headers: 16-bit types are opaque 2-byte storage on the
//    host, arithmetic happens in the sm_100a kernels behind the C ABI (include/dlnb.h).
//  * `_Stream` + CREATE/SYNC/DESTROY_STREAM (reference :91-130) map onto slots of the
//    communication stream owned by libdlnb; they are kept as no-cost shims.
#ifndef DATA_TYPES_HPP
#define DATA_TYPES_HPP

#include <dlnb.h>
#include <stdint.h>

#define PROXY_ENABLE_B200 1

struct proxy_half { uint16_t bits; };      // IEEE binary16 storage
struct proxy_bfloat16 { uint16_t bits; };  // bfloat16 storage

template <typename T> struct ProxyDtype;
template <> struct ProxyDtype<float> { static constexpr int value = DLNB_F32; static constexpr const char *name = "fp32"; };
template <> struct ProxyDtype<proxy_half> { static constexpr int value = DLNB_F16; static constexpr const char *name = "fp16"; };
template <> struct ProxyDtype<proxy_bfloat16> { static constexpr int value = DLNB_BF16; static constexpr const char *name = "bf16"; };

#if defined(PROXY_FLOAT32)
using _FLOAT = float;
#elif defined(PROXY_BFLOAT16)
using _FLOAT = proxy_bfloat16;
#else
using _FLOAT = proxy_half;
#endif
#define B200_FLOAT_TYPE (ProxyDtype<_FLOAT>::value)

// Communicator handle type (reference: Proxy_CommType = ncclComm_t / MPI_Comm, :82-88)
using Proxy_CommType = dlnb_comm_t;

// Streams: a "stream" of the reference is a slot index here.
using _Stream = int;
#define CREATE_STREAM(stream) ((stream) = 0)
#define DESTROY_STREAM(stream) ((void)(stream))
#define SYNC_STREAM(stream) ((void)(stream))

#endif // DATA_TYPES_HPP
