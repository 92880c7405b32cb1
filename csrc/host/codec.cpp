// Byte-exact codecs for the reference's two binary formats:
//  * tensor record of model snapshots (include/tensor/tensor.hpp:625-653): size_t shape[4] + raw float32 data;
//  * pipeline wire messages (include/pipeline/binary_serializer.hpp:27-176, message.hpp:23-37):
//      FixedHeader{u8 version=1, u8 endianness (0 little, 1 big), u64 body length}
//      body = u64 len + recipient, u64 len + sender, u16 command, u64 payload_type, payload
//      payload_type: 0 none | 1 Job{u64 micro_batch_id, tensor} | 2 string{u64 len, bytes} | 3 bool{u8}
//      tensor = u64 ndims, u64 dims[ndims], raw float32
// Only control-plane traffic and checkpoints use these; activations never leave the GPUs.
#include <string.h>

#include "host_common.h"

namespace {
struct Writer {
  uint8_t* p;
  size_t cap, off;
  bool ok;
  void put(const void* src, size_t n) {
    if (p && off + n <= cap) memcpy(p + off, src, n);
    else if (p) ok = false;
    off += n;
  }
  template <class T> void val(T v) { put(&v, sizeof(T)); }
};
struct Reader {
  const uint8_t* p;
  size_t size, off;
  bool ok;
  bool get(void* dst, size_t n) {
    if (off + n > size) { ok = false; return false; }
    memcpy(dst, p + off, n);
    off += n;
    return true;
  }
  template <class T> T val() { T v{}; get(&v, sizeof(T)); return v; }
};
}  // namespace

DCNN_HOST_API int dcnn_host_endianness() {
  const uint32_t probe = 0x01020304;
  return *reinterpret_cast<const uint8_t*>(&probe) == 1 ? 1 : 0;
}

// Returns the encoded size.  With out == NULL only the size is computed.  -1 if `cap` is too small.
DCNN_HOST_API long long dcnn_encode_message(uint8_t* out, size_t cap, const char* recipient, const char* sender,
                                            uint16_t command, uint64_t payload_type, uint64_t micro_batch_id,
                                            const uint64_t* dims, uint64_t ndims, const float* tensor,
                                            const uint8_t* str, uint64_t str_len, uint8_t flag) {
  Writer w{out, cap, 0, true};
  // body first (its length goes into the fixed header)
  Writer body{nullptr, 0, 0, true};
  for (int pass = 0; pass < 2; ++pass) {
    Writer& b = pass == 0 ? body : w;
    if (pass == 1) {
      w.val<uint8_t>(1);
      w.val<uint8_t>((uint8_t)dcnn_host_endianness());
      w.val<uint64_t>(body.off);
    }
    const uint64_t rl = strlen(recipient), sl = strlen(sender);
    b.val<uint64_t>(rl); b.put(recipient, rl);
    b.val<uint64_t>(sl); b.put(sender, sl);
    b.val<uint16_t>(command);
    b.val<uint64_t>(payload_type);
    if (payload_type == 1) {
      b.val<uint64_t>(micro_batch_id);
      b.val<uint64_t>(ndims);
      uint64_t numel = 1;
      for (uint64_t i = 0; i < ndims; ++i) { b.val<uint64_t>(dims[i]); numel *= dims[i]; }
      if (ndims == 0) numel = 0;
      b.put(tensor, numel * sizeof(float));
    } else if (payload_type == 2) {
      b.val<uint64_t>(str_len);
      b.put(str, str_len);
    } else if (payload_type == 3) {
      b.val<uint8_t>(flag ? 1 : 0);
    }
  }
  return w.ok ? (long long)w.off : -1;
}

// Parses the fixed header + message header + payload descriptor.  Offsets point into `buf`.
// Returns 0 on success, -1 on truncated/invalid input.
DCNN_HOST_API int dcnn_decode_message(const uint8_t* buf, size_t size, uint64_t* body_len, uint64_t* recipient_off,
                                      uint64_t* recipient_len, uint64_t* sender_off, uint64_t* sender_len,
                                      uint16_t* command, uint64_t* payload_type, uint64_t* micro_batch_id,
                                      uint64_t* dims /* >= 8 */, uint64_t* ndims, uint64_t* data_off,
                                      uint64_t* data_len, uint8_t* flag) {
  Reader r{buf, size, 0, true};
  const uint8_t version = r.val<uint8_t>();
  const uint8_t endian = r.val<uint8_t>();
  *body_len = r.val<uint64_t>();
  if (!r.ok || version != 1 || endian != (uint8_t)dcnn_host_endianness()) return -1;  // foreign endianness unsupported
  *recipient_len = r.val<uint64_t>(); *recipient_off = r.off; r.off += *recipient_len;
  *sender_len = r.val<uint64_t>();    *sender_off = r.off;    r.off += *sender_len;
  if (r.off > size) return -1;
  *command = r.val<uint16_t>();
  *payload_type = r.val<uint64_t>();
  *ndims = 0; *data_off = 0; *data_len = 0; *flag = 0; *micro_batch_id = 0;
  if (*payload_type == 1) {
    *micro_batch_id = r.val<uint64_t>();
    *ndims = r.val<uint64_t>();
    if (*ndims > 8) return -1;
    uint64_t numel = *ndims ? 1 : 0;
    for (uint64_t i = 0; i < *ndims; ++i) { dims[i] = r.val<uint64_t>(); numel *= dims[i]; }
    *data_off = r.off; *data_len = numel * sizeof(float); r.off += *data_len;
  } else if (*payload_type == 2) {
    *data_len = r.val<uint64_t>(); *data_off = r.off; r.off += *data_len;
  } else if (*payload_type == 3) {
    *flag = r.val<uint8_t>();
  } else if (*payload_type != 0) {
    return -1;
  }
  return (r.ok && r.off <= size) ? 0 : -1;
}

// Tensor record: 32-byte header (4 x size_t) + data.  Returns bytes written / needed.
DCNN_HOST_API long long dcnn_encode_tensor_record(uint8_t* out, size_t cap, const uint64_t shape4[4],
                                                  const float* data) {
  const uint64_t numel = shape4[0] * shape4[1] * shape4[2] * shape4[3];
  const size_t need = 32 + numel * sizeof(float);
  if (!out) return (long long)need;
  if (cap < need) return -1;
  memcpy(out, shape4, 32);
  memcpy(out + 32, data, numel * sizeof(float));
  return (long long)need;
}
