/* tcb200.h - C ABI of libtcb200.so: the B200 (sm_100a) tensor-marshalling path of the Triton
 * client, as a drop-in behind tritonclient's own Python API.
 *
 * Style follows the reference's only C shim, libcshm
 * (/root/reference/src/python/library/tritonclient/utils/shared_memory/shared_memory.h:34-53):
 * every export returns int (0 = success, negative = error class), results come back through
 * pointer parameters, handles are opaque void*, and the Python side binds it with ctypes
 * (shared_memory/__init__.py:48-72). No torch / numpy / C++ types cross this boundary.
 *
 * All `stream` parameters are a cudaStream_t passed as void* (NULL = the legacy default
 * stream). Unless a function says otherwise it only ENQUEUES work on that stream and
 * returns; pointers named *_dev must be device-accessible on `device_id`.
 */
#ifndef TCB200_H_
#define TCB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error classes (libcshm uses -2..-6 the same way: shared_memory/__init__.py:314-340) ---- */
#define TCB_OK 0
#define TCB_ERR_INVALID_ARG (-1)  /* bad argument (null pointer, unknown enum, negative size)   */
#define TCB_ERR_CUDA (-2)         /* a CUDA runtime call failed; TcbLastError() has its text    */
#define TCB_ERR_OUT_OF_RANGE (-3) /* offset + size exceeds the region / output capacity        */
#define TCB_ERR_TRUNCATED (-4)    /* BYTES stream ends inside a length prefix or a payload      */
#define TCB_ERR_UNSUPPORTED (-5)  /* valid request this build does not handle                   */
#define TCB_ERR_NOMEM (-6)        /* host or device allocation failed                           */
#define TCB_ERR_INTERNAL (-7)     /* device-side invariant violated (should never happen)       */

/* Text of the last error raised on the calling thread (thread-local, never NULL). */
const char* TcbLastError(void);
/* Library version / build info, e.g. "tcb200 0.1 sm_100a". */
const char* TcbVersion(void);
/* Number of kernels this library has launched in this process (for bench.py's gpu_launches). */
uint64_t TcbKernelLaunchCount(void);

/* =============================================================================================
 * CUDA shared-memory regions.
 * Replaces the cuda-python calls of
 *   tritonclient/utils/cuda_shared_memory/__init__.py:107-149  create_shared_memory_region
 *   tritonclient/utils/cuda_shared_memory/__init__.py:152-170  get_raw_handle
 *   tritonclient/utils/cuda_shared_memory/__init__.py:173-239  set_shared_memory_region
 *   tritonclient/utils/cuda_shared_memory/__init__.py:242-325  get_contents_as_numpy (the D2H)
 *   tritonclient/utils/cuda_shared_memory/_utils.py:88-100     CudaSharedMemoryRegion.__del__
 * A region is a plain cudaMalloc allocation (so that its cudaIpcMemHandle_t has no offset and
 * can be registered with the server as is) plus its IPC handle.
 * ============================================================================================= */

/* cudaSetDevice(device_id) + cudaMalloc(byte_size) + cudaIpcGetMemHandle; restores the previous
 * device. byte_size 0 is allowed (the reference lets cudaMalloc decide; here it yields a NULL
 * base). */
int TcbRegionCreate(const char* triton_shm_name, uint64_t byte_size, int device_id, void** region);

/* base_addr: device pointer as an integer; ipc_handle: the 64 raw bytes of cudaIpcMemHandle_t
 * (what the reference base64-encodes from `.reserved`). Any out pointer may be NULL. */
int TcbRegionInfo(void* region, uint64_t* base_addr, uint64_t* byte_size, int* device_id,
                  unsigned char ipc_handle[64]);

/* Copy byte_size bytes from src (host or device pointer, cudaMemcpyDefault) to region+offset on
 * `stream`; bounds-checked (the reference is not). Does NOT synchronise: call TcbStreamSync, as
 * set_shared_memory_region does once after all arrays (cuda_shared_memory/__init__.py:231). */
int TcbRegionSet(void* region, uint64_t offset, const void* src, uint64_t byte_size, void* stream);

/* Copy byte_size bytes from region+offset to dst (host or device) on `stream`, then synchronise
 * the stream (cuda_shared_memory/__init__.py:266-276; the reference always copies the whole
 * region, this copies what is asked). */
int TcbRegionGet(void* region, uint64_t offset, void* dst, uint64_t byte_size, void* stream);

/* cudaFree on the region's device (restoring the previous one) and release the handle. */
int TcbRegionDestroy(void* region);

/* One stream per device, process-global, like _get_or_create_global_cuda_stream
 * (cuda_shared_memory/__init__.py:66-70). Never destroyed before process exit. */
int TcbGlobalStream(int device_id, void** stream);
int TcbStreamSync(int device_id, void* stream);

/* Pinned host staging (used by the Python layer for H2D/D2H at the PCIe edge). */
int TcbHostAlloc(uint64_t byte_size, void** ptr);
int TcbHostFree(void* ptr);
/* Plain device scratch/workspace allocation for callers without torch. */
int TcbDeviceAlloc(int device_id, uint64_t byte_size, void** ptr_dev);
int TcbDeviceFree(int device_id, void* ptr_dev);
/* cudaMemcpyAsync(cudaMemcpyDefault) + optional sync. */
int TcbMemcpy(int device_id, void* dst, const void* src, uint64_t byte_size, void* stream, int sync);

/* =============================================================================================
 * System (POSIX) shared-memory regions.
 * Replaces libcshm: tritonclient/utils/shared_memory/shared_memory.cc:76-105 SharedMemoryRegionCreate,
 * :107-116 SharedMemoryRegionSet, :118-131 GetSharedMemoryHandleInfo, :133-148
 * SharedMemoryRegionDestroy (declared in shared_memory.h:34-53). Same return codes for the failures
 * libcshm knows (-2 shm_open, -3 ftruncate, -4 mmap, -5 shm_unlink, -6 munmap).
 * pin != 0 page-locks the mapping (cudaHostRegister, best effort) so that device <-> region
 * copies run by DMA; copies are bounds-checked (libcshm's are not).
 * ============================================================================================= */
int TcbShmRegionCreate(const char* triton_shm_name, const char* shm_key, uint64_t byte_size,
                       int pin, void** shm_handle);
int TcbShmRegionSet(void* shm_handle, uint64_t offset, uint64_t byte_size, const void* data);
/* Any out pointer may be NULL. offset is always 0 (libcshm maps the whole object too). */
int TcbShmRegionInfo(void* shm_handle, void** shm_addr, const char** shm_key, int* shm_fd,
                     uint64_t* offset, uint64_t* byte_size, int* pinned);
/* Copy byte_size bytes between device memory and region + offset on `stream` (to_region != 0:
 * device -> region), then synchronise the stream. No counterpart in libcshm: the reference moves
 * device results to a system region through a numpy array. */
int TcbShmRegionCopyDevice(void* shm_handle, uint64_t offset, void* dev_ptr, uint64_t byte_size,
                           int device_id, int to_region, void* stream);
int TcbShmRegionDestroy(void* shm_handle);

/* =============================================================================================
 * Image preprocess.
 * Replaces image_client.py:154-193 `preprocess` (PIL convert + resize(BILINEAR) + numpy
 * astype/scale/transpose, /root/reference/src/python/examples/image_client.py) and the
 * np.stack + H2D that follow it (image_client.py:452; cuda_shared_memory/__init__.py:222-229):
 * a batch of uint8 images already on the device is resampled with Pillow's exact 8-bit
 * arithmetic, mapped through a 256-entry table per channel and written in NHWC or NCHW order
 * straight to `dst_dev` (normally region base + offset).
 * ============================================================================================= */

#define TCB_LAYOUT_NHWC 1 /* ModelInput.FORMAT_NHWC */
#define TCB_LAYOUT_NCHW 2 /* ModelInput.FORMAT_NCHW */

typedef struct TcbImage {
  uint64_t src_dev;     /* device pointer to the first pixel, row-major, interleaved channels */
  int32_t width;        /* pixels */
  int32_t height;       /* rows */
  int32_t pitch;        /* bytes between rows (>= width * src_channels) */
  int32_t src_channels; /* 1 = L, 3 = RGB, 4 = RGBA (alpha ignored) */
} TcbImage;

/* A plan fixes the geometry of a batch (per-image source sizes, target c/h/w, layout, value
 * table) and owns everything derived from it on the device: Pillow's fixed-point coefficient
 * tables, per-image descriptors, and scratch. Build once, run every step.
 *
 *  c        : 1 (convert("L")) or 3 (convert("RGB"))                      image_client.py:161-164
 *  out_h/w  : target size                                                 image_client.py:166
 *  layout   : TCB_LAYOUT_NHWC | TCB_LAYOUT_NCHW                           image_client.py:185-188
 *  lut      : host pointer, c * 256 entries of elem_size bytes: lut[ch][v] is the output value
 *             for resized pixel value v in channel ch. The Python layer fills it by running
 *             the reference's own numpy expressions (astype / INCEPTION / VGG,
 *             image_client.py:171-182) on arange(256), which makes the dtype conversion and
 *             scaling bit-exact for every dtype, including numpy's fp16 double rounding.
 *  elem_size: 1, 2, 4 or 8
 *  path     : 0 = auto (a fused kernel where one applies, two-pass kernels otherwise), 1 = force the
 *             two-pass kernels, 2 = require a fused kernel (TCB_ERR_UNSUPPORTED if any image's
 *             geometry does not fit one), 3 = like 2 but never the IMMA variant. There are two fused
 *             kernels: an IMMA (int8 tensor-core) one for RGB sources whose rows are 16-byte aligned
 *             (base pointer and pitch multiples of 16) and a dp4a one for everything else.
 * Image i is written at dst_dev + i * c*out_h*out_w*elem_size. */
int TcbPreprocessPlanCreate(int device_id, const TcbImage* images, int n_images, int c, int out_h,
                            int out_w, int layout, const void* lut, int elem_size, int path,
                            void** plan);
/* Source pointers/pitches may change between runs (same sizes); pass NULL to keep the ones
 * given at creation (src_dev may be 0 at creation if every run supplies pointers). The launch
 * (which kernel, how much shared memory) is fixed at creation: a run whose new pointers or pitches
 * no longer fit it - rows that lose the alignment the chosen kernel needs, a pitch far larger than
 * the plan's - returns TCB_ERR_UNSUPPORTED and changes nothing; create a new plan then. A run that
 * fails validation leaves the plan as it was.
 * A plan is NOT re-entrant: it owns one set of device descriptors, rewritten on `stream` when the
 * pointers change, so run one plan from one thread / one stream at a time (different plans are
 * independent; there is no process-global state). The fused kernels fetch source rows with 16-byte
 * granular bulk copies: the dp4a kernel may read up to 15 bytes before the first and after the last
 * byte of an image whose rows are not 16-byte aligned (never outside the 16-byte blocks the image
 * touches).
 * dst_capacity is the number of bytes available at dst_dev. */
int TcbPreprocessPlanRun(void* plan, const TcbImage* images_or_null, void* dst_dev,
                         uint64_t dst_capacity, void* stream);
/* Bytes of plan-owned scratch on the device (the two-pass path's uint8 intermediate). */
int TcbPreprocessPlanInfo(void* plan, uint64_t* scratch_bytes, uint64_t* out_bytes_per_image,
                          int* n_fused_images);
int TcbPreprocessPlanDestroy(void* plan);

/* Pillow's coefficient set-up for one axis, exposed for tests: bounds[out*2] = (first tap,
 * tap count), kk[out*ksize] 22-bit fixed-point weights. Returns ksize (> 0) or an error. Host
 * only, no GPU needed. */
int TcbResampleCoeffs(int in_size, int out_size, int32_t* bounds, int32_t* kk, int kk_capacity);

/* =============================================================================================
 * BYTES tensors.
 * Replaces tritonclient/utils/__init__.py:193-246 serialize_byte_tensor (np.nditer +
 * struct.pack loop + b"".join) and :249-276 deserialize_bytes_tensor / the inlined copy at
 * cuda_shared_memory/__init__.py:305-323 (unpack_from loop).
 * Wire format: per element, in C order: uint32 little-endian length, then the payload.
 * ============================================================================================= */

/* Workspace (device) needed by the calls below for n elements / nbytes of stream. */
uint64_t TcbBytesSerializeWorkspace(uint64_t n_elements);
uint64_t TcbBytesDeserializeWorkspace(uint64_t stream_bytes);

/* payload_dev : all payloads back to back in element order (P bytes)
 * lengths_dev : n lengths; length_size 4 (int32/uint32) or 8 (int64)
 * dst_dev     : receives P + 4n bytes; dst_capacity checked against the true total
 * total_dev   : optional device uint64 that receives P + 4n (may be NULL)
 * total_host  : optional host pointer; when non-NULL the call synchronises the stream and
 *               stores P + 4n (or returns TCB_ERR_OUT_OF_RANGE if it exceeded dst_capacity, in
 *               which case nothing past the capacity was written). */
int TcbBytesSerialize(int device_id, const void* payload_dev, const void* lengths_dev,
                      int length_size, uint64_t n_elements, void* dst_dev, uint64_t dst_capacity,
                      void* workspace_dev, uint64_t workspace_bytes, uint64_t* total_dev,
                      uint64_t* total_host, void* stream);

/* Same result as TcbBytesSerialize, for callers that must not allocate. No counterpart in the
 * reference (its serialiser is a host-side Python loop; there is nothing to bind this to).
 *
 * Who needs it: code that runs while the device's memory is already spoken for or while an
 * allocation would stall it - a serving process that handed its whole HBM budget to its own
 * arena, a request handler on a latency budget (cudaMalloc takes a driver-wide lock and
 * synchronises the device), a callback or library hook that is simply not allowed to call the
 * allocator. TcbBytesSerialize asks such a caller for a workspace it cannot create.
 *
 * Here the scan scratch is a fixed array inside the library image instead (131 144 bytes of
 * device memory, present on every GPU that loads libtcb200.so for the life of the process), so
 * the call takes only inputs and outputs and calls no cudaMalloc / cudaFree / allocating
 * library routine.
 *   - Limit: n_elements <= 16 777 216 per call (one status word per 1024 elements); more is
 *     refused with TCB_ERR_OUT_OF_RANGE and a message that says so.
 *   - The first call on each device looks up the array's device address (a one-off set-up
 *     step under that device); make that first call before starting any stream capture.
 *   - All calls on a device share the one array. Each call clears the part it uses on `stream`
 *     before its kernel runs, so back-to-back calls on one stream are fine; two calls whose GPU
 *     work can overlap (different streams, different threads) race on it and must be ordered by
 *     the caller.
 *   - Errors found on the device (destination too small: TCB_ERR_CAPACITY) are reported through
 *     the synchronising form only, i.e. when total_host is given. With total_host == NULL the call
 *     returns once the work is enqueued and has nowhere to report them (there is no total_dev
 *     here): size dst_capacity from the lengths (payload bytes + 4 per element) before using it.
 * Arguments as TcbBytesSerialize, minus the workspace and total_dev. */
int TcbBytesSerializeNoAlloc(int device_id, const void* payload_dev, const void* lengths_dev,
                             int length_size, uint64_t n_elements, void* dst_dev,
                             uint64_t dst_capacity, uint64_t* total_host, void* stream);

/* Fixed-width rows (numpy `S<k>` arrays: n rows of `width` bytes, element = row with trailing
 * NULs stripped, utils/__init__.py:239). Same outputs as TcbBytesSerialize. */
int TcbBytesSerializeFixedWidth(int device_id, const void* rows_dev, uint32_t width,
                                uint64_t n_elements, void* dst_dev, uint64_t dst_capacity,
                                void* workspace_dev, uint64_t workspace_bytes,
                                uint64_t* total_dev, uint64_t* total_host, void* stream);

/* stream_dev     : the serialized bytes (stream_bytes of them), e.g. region base
 * max_elements   : < 0 -> decode to the end of the buffer (deserialize_bytes_tensor);
 *                  >= 0 -> stop after that many elements (get_contents_as_numpy)
 * offsets_dev    : int64[offsets_capacity + 1]; element i is payload[offsets[i]:offsets[i+1]]
 * payload_dev    : receives the payloads back to back (capacity payload_capacity bytes)
 * n_elements_host, payload_bytes_host : results (host). The call synchronises the stream.
 * Returns TCB_ERR_TRUNCATED where the reference raises struct.error.
 * A stream of at most 16 KiB (a request's handful of strings) is decoded by one single-CTA kernel
 * launch and one host round trip instead of the seven-launch chunked path. For the BYTES entry
 * points "device" pointers may also be page-locked host memory (cudaHostAlloc / cudaHostRegister:
 * device-addressable under unified addressing): the Python layer serialises small tensors straight
 * out of and into one pinned staging buffer that way. Keep a stream to be decoded in device memory
 * though - the decoder follows a chain of dependent loads, ~1.5 us per element over PCIe. */
int TcbBytesDeserialize(int device_id, const void* stream_dev, uint64_t stream_bytes,
                        int64_t max_elements, int64_t* offsets_dev, uint64_t offsets_capacity,
                        void* payload_dev, uint64_t payload_capacity, void* workspace_dev,
                        uint64_t workspace_bytes, uint64_t* n_elements_host,
                        uint64_t* payload_bytes_host, void* stream);

/* The same decode in two calls, for callers that want to size their buffers exactly (the combined
 * call needs offsets for the worst case of stream_bytes / 4 elements - 4 GB for a 2 GB stream).
 * Both replace the same reference loop as TcbBytesDeserialize (utils/__init__.py:249-276).
 *   Count: chunked offset chase + verification only. *n_elements_host = elements that will be decoded,
 *          *payload_bound_host = stream_bytes - 4 * n (an upper bound of the payload bytes, exact when
 *          the whole buffer is decoded). Returns TCB_ERR_TRUNCATED like the combined call.
 *   Fill : writes offsets (n + 1 entries) and the payload. Must follow a successful Count on the SAME
 *          workspace, stream_dev, stream_bytes and max_elements, with nothing else having used the
 *          workspace in between. Both calls synchronise the stream. */
int TcbBytesDeserializeCount(int device_id, const void* stream_dev, uint64_t stream_bytes,
                             int64_t max_elements, void* workspace_dev, uint64_t workspace_bytes,
                             uint64_t* n_elements_host, uint64_t* payload_bound_host, void* stream);
int TcbBytesDeserializeFill(int device_id, const void* stream_dev, uint64_t stream_bytes,
                            int64_t max_elements, int64_t* offsets_dev, uint64_t offsets_capacity,
                            void* payload_dev, uint64_t payload_capacity, void* workspace_dev,
                            uint64_t workspace_bytes, uint64_t* n_elements_host,
                            uint64_t* payload_bytes_host, void* stream);

/* =============================================================================================
 * BF16 tensors.
 * Replaces tritonclient/utils/__init__.py:279-323 serialize_bf16_tensor (np.nditer +
 * struct.pack("<f", x)[2:4] per element + b"".join) and :326-348 deserialize_bf16_tensor
 * (unpack_from("<2s") + struct.unpack("<f", b"\x00\x00" + sb) per element).
 * Wire format: per element, in C order, the upper 16 bits of the float32 bit pattern, little
 * endian - plain truncation, no rounding. One quirk is kept: the reference goes through Python
 * floats, which turns a signalling NaN into a quiet one (mantissa bit 22 set) in both directions.
 * ============================================================================================= */

/* src_f32_dev: n_elements float32 (4-byte aligned); dst_dev receives 2 * n_elements bytes
 * (2-byte aligned; 128-bit stores are used when source and destination alignments agree). */
int TcbBf16Serialize(int device_id, const void* src_f32_dev, uint64_t n_elements, void* dst_dev,
                     uint64_t dst_capacity, void* stream);
/* src_dev: src_bytes of bfloat16 wire data; dst_f32_dev receives src_bytes / 2 float32 values
 * (low 16 bits zero). An odd src_bytes returns TCB_ERR_TRUNCATED where the reference's
 * unpack_from raises struct.error. */
int TcbBf16Deserialize(int device_id, const void* src_dev, uint64_t src_bytes, void* dst_f32_dev,
                       uint64_t dst_capacity, void* stream);

/* =============================================================================================
 * Classification strings: field split on the device (postprocess half)
 * Replaces the per-character join and ':' split of every element in
 *   /root/reference/src/python/examples/image_client.py:196-216 (postprocess)
 * for outputs decoded by TcbBytesDeserialize(Fill): payload_dev + offsets_dev (num_elements + 1
 * int64, as written by the deserialiser). fields_dev receives 4 int64 per element: the positions,
 * relative to the element's first byte, of its first, second and third ':' (-1 where it has fewer),
 * and the element's length. The caller slices cls[0] = [0, f0), cls[1] = (f0, f1), cls[2] =
 * (f1, f2 or end) - what str.split(":")[0:3] yields; an element with f1 == -1 makes the reference
 * raise IndexError.
 * ============================================================================================= */
int TcbClassificationFields(int device_id, const void* payload_dev, const int64_t* offsets_dev,
                            int64_t num_elements, int64_t* fields_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TCB200_H_ */
