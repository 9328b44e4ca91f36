// Test harness (not shipped): the device field parser of csrc/text_parse.inl compiled for the
// host, so that tests/test_text_parse_host.py can fuzz it against the C library's sscanf("%f")
// without a GPU.  Built by the test into tests/native/_build/.
#include "../../hadoop-1.0.3-gpu_b200/csrc/text_parse.inl"

extern "C" int text_parse_host(const char *s, int len, uint32_t *bits, int *trouble) {
  bool t = false;
  const bool ok = hgm_text::parse_field(reinterpret_cast<const unsigned char *>(s), len, bits, &t);
  *trouble = t ? 1 : 0;
  return ok ? 0 : -1;
}

// many fields at once: offsets[i]..offsets[i+1] delimit field i; rc[i] = 0 ok, -1 rejected, 1 trouble
extern "C" void text_parse_host_many(const char *buf, const long long *offsets, long long n, uint32_t *bits, int *rc) {
  for (long long i = 0; i < n; ++i) {
    bool t = false;
    const bool ok = hgm_text::parse_field(reinterpret_cast<const unsigned char *>(buf + offsets[i]),
                                          (int)(offsets[i + 1] - offsets[i]), &bits[i], &t);
    rc[i] = !ok ? -1 : (t ? 1 : 0);
  }
}
