// Library-wide C-ABI entry points: version and last-error string.
#include "tdl_common.cuh"

TDL_API const char* tdl_last_error() { return tdl::err_buf(); }

// Encodes the ABI revision; bump when an existing signature changes.
TDL_API int tdl_version() { return 101; }  // 101: tdl_bn_apply gained mask_out, tdl_bn_bwd_* gained z_is_bitmask

// Number of SMs of the current device as the kernels' launch heuristics see it.
TDL_API int tdl_num_sms() { return tdl::num_sms(); }

// CRC-32C (Castagnoli), slicing-by-8 on the host: the checksum TensorFlow's tensor-bundle checkpoints use for both
// the index table blocks and the tensor payloads (estimator/tf_bundle.py).  `crc` is the running value (0 to start).
static uint32_t g_crc32c_table[8][256];
static bool g_crc32c_ready = false;
static void crc32c_init() {
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : (c >> 1);
    g_crc32c_table[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i)
    for (int t = 1; t < 8; ++t)
      g_crc32c_table[t][i] = (g_crc32c_table[t - 1][i] >> 8) ^ g_crc32c_table[0][g_crc32c_table[t - 1][i] & 0xff];
  g_crc32c_ready = true;
}
TDL_API uint32_t tdl_crc32c(const void* data, uint64_t n, uint32_t crc) {
  if (!g_crc32c_ready) crc32c_init();
  const unsigned char* p = static_cast<const unsigned char*>(data);
  uint32_t c = ~crc;
  while (n >= 8) {
    uint32_t lo, hi;
    memcpy(&lo, p, 4);
    memcpy(&hi, p + 4, 4);
    lo ^= c;
    c = g_crc32c_table[7][lo & 0xff] ^ g_crc32c_table[6][(lo >> 8) & 0xff] ^ g_crc32c_table[5][(lo >> 16) & 0xff] ^
        g_crc32c_table[4][lo >> 24] ^ g_crc32c_table[3][hi & 0xff] ^ g_crc32c_table[2][(hi >> 8) & 0xff] ^
        g_crc32c_table[1][(hi >> 16) & 0xff] ^ g_crc32c_table[0][hi >> 24];
    p += 8;
    n -= 8;
  }
  while (n--) c = (c >> 8) ^ g_crc32c_table[0][(c ^ *p++) & 0xff];
  return ~c;
}
