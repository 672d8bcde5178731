// Native reader of the framework's weight files (format: cuda-mpi-gpu-cluster-programming_b200/utils/checkpoint.py:
// 16-byte preamble, JSON header, raw little-endian float32 tensors 256 B aligned in a data section that starts at
// the next multiple of 4096).  A native host (the V1 / V2 binaries, a C++ serving process on the C ABI) loads the
// same files the Python side writes.  The header is the JSON this project generates itself, so the scanner below
// only understands that shape: an object per tensor with string "name", integer "offset" / "nbytes", a "shape" list
// and a "fnv1a64" string; it does not try to be a JSON parser.
#include "b200cnn.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace {

bool read_header(FILE* f, std::string* json, long* data_start) {
    unsigned char pre[16];
    if (fread(pre, 1, 16, f) != 16 || memcmp(pre, "B200CNNW", 8) != 0) return false;
    const uint32_t ver = pre[8] | (pre[9] << 8) | (pre[10] << 16) | ((uint32_t)pre[11] << 24);
    const uint32_t n = pre[12] | (pre[13] << 8) | (pre[14] << 16) | ((uint32_t)pre[15] << 24);
    if (ver != 1 || n == 0 || n > (64u << 20)) return false;
    json->resize(n);
    if (fread(&(*json)[0], 1, n, f) != n) return false;
    *data_start = (long)((16 + (size_t)n + 4095) / 4096 * 4096);
    return true;
}

// the {...} of the tensor whose "name" is `name`
bool find_entry(const std::string& js, const char* name, std::string* entry) {
    const std::string key = std::string("\"name\": \"") + name + "\"";
    const size_t at = js.find(key);
    if (at == std::string::npos) return false;
    const size_t lo = js.rfind('{', at), hi = js.find('}', at);
    if (lo == std::string::npos || hi == std::string::npos) return false;
    *entry = js.substr(lo, hi - lo + 1);
    return true;
}

bool int_field(const std::string& e, const char* field, long long* out) {
    const std::string key = std::string("\"") + field + "\": ";
    const size_t at = e.find(key);
    if (at == std::string::npos) return false;
    char* end = nullptr;
    *out = strtoll(e.c_str() + at + key.size(), &end, 10);
    return end != e.c_str() + at + key.size();
}

}  // namespace

extern "C" int b200_weights_read(const char* path, const char* name, float* dst, size_t capacity, size_t* count,
                                 int* shape, int* ndim) {
    if (!path || !name || !count) return B200_ERR_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return B200_ERR_ARG;
    std::string js, e;
    long base = 0;
    int rc = B200_OK;
    long long off = 0, nbytes = 0;
    if (!read_header(f, &js, &base) || !find_entry(js, name, &e) || !int_field(e, "offset", &off) ||
        !int_field(e, "nbytes", &nbytes) || off < 0 || nbytes < 0 || nbytes % 4 != 0) {
        fclose(f);
        return B200_ERR_UNSUPPORTED;
    }
    *count = (size_t)nbytes / 4;
    if (shape && ndim) {
        *ndim = 0;
        const size_t s = e.find("\"shape\": [");
        if (s != std::string::npos) {
            const char* p = e.c_str() + s + 10;
            while (*p && *p != ']' && *ndim < 4) {
                char* end = nullptr;
                const long v = strtol(p, &end, 10);
                if (end == p) break;
                shape[(*ndim)++] = (int)v;
                p = end;
                while (*p == ',' || *p == ' ') ++p;
            }
        }
    }
    if (dst) {
        if (capacity < *count) rc = B200_ERR_RANGE;
        else if (fseek(f, base + (long)off, SEEK_SET) != 0 || fread(dst, 4, *count, f) != *count) rc = B200_ERR_RANGE;
        else {
            const size_t h = e.find("\"fnv1a64\": \"0x");
            if (h != std::string::npos) {
                const unsigned long long want = strtoull(e.c_str() + h + 14, nullptr, 16);
                if (want != (unsigned long long)b200_fnv1a64(dst, (size_t)nbytes)) rc = B200_ERR_RANGE;   // corrupt
            }
        }
    }
    fclose(f);
    return rc;
}
