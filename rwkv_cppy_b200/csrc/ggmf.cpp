// ggmf container reader: the whole file is mapped read-only and tensors are views into
// the mapping (the reference reads tensor by tensor with fread, rwkv_file_format.inc:240-313;
// the checks and the error flags they raise follow rwkv_file_format.inc:115-197 and
// rwkv_model_loading.inc:295-317).
#include "common.h"

#include <cstdarg>
#include <cstring>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace rb {

thread_local int  g_last_error   = RWKV_ERROR_NONE;
thread_local bool g_print_errors = true;

void global_fail(int flags, const char * file, int line, const char * fmt, ...) {
    g_last_error |= flags;
    if (!g_print_errors) return;
    va_list ap;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fprintf(stderr, "\n%s:%d: error 0x%x\n", file, line, flags);
}

static const char * const k_type_names[T_COUNT] = {
    "FP32", "FP16", "Q4_0", "Q4_1", "Q4_1_O", "Q4_2", "Q4_3", "Q5_0", "Q5_1", "Q8_0"
};

const char * type_name(uint32_t t) { return t < T_COUNT ? k_type_names[t] : "unknown"; }

int type_from_name(const char * s) {
    for (int i = 0; i < (int) T_COUNT; i++) if (strcmp(s, k_type_names[i]) == 0) return i;
    return -1;
}

GgmfFile::~GgmfFile() {
    if (map_base && map_base != MAP_FAILED) munmap(map_base, map_size);
    if (fd >= 0) close(fd);
}

namespace {
struct Cursor {
    const uint8_t * p; size_t left;
    bool u32(uint32_t & v) { if (left < 4) return false; memcpy(&v, p, 4); p += 4; left -= 4; return true; }
    bool skip(size_t n) { if (left < n) return false; p += n; left -= n; return true; }
};
}

bool ggmf_open(const char * path, GgmfFile & f) {
    f.fd = open(path, O_RDONLY);
    RB_CHECK(f.fd >= 0, RWKV_ERROR_FILE | RWKV_ERROR_FILE_OPEN, false, "Failed to open file %s", path);
    struct stat st;
    RB_CHECK(fstat(f.fd, &st) == 0, RWKV_ERROR_FILE | RWKV_ERROR_FILE_STAT, false, "Failed to stat file %s", path);
    f.map_size = (size_t) st.st_size;
    RB_CHECK(f.map_size >= 24, RWKV_ERROR_FILE | RWKV_ERROR_FILE_READ, false, "Invalid file header");
    f.map_base = mmap(nullptr, f.map_size, PROT_READ, MAP_PRIVATE, f.fd, 0);
    RB_CHECK(f.map_base != MAP_FAILED, RWKV_ERROR_FILE | RWKV_ERROR_FILE_READ, false, "Failed to map file %s", path);
    madvise(f.map_base, f.map_size, MADV_SEQUENTIAL);

    Cursor c{(const uint8_t *) f.map_base, f.map_size};
    uint32_t magic = 0;
    c.u32(magic); c.u32(f.version); c.u32(f.n_vocab); c.u32(f.n_embed); c.u32(f.n_layer); c.u32(f.data_type);
    RB_CHECK(magic == RWKV_FILE_MAGIC, RWKV_ERROR_FILE | RWKV_ERROR_FILE_MAGIC, false, "Invalid file header");
    RB_CHECK(f.version >= RWKV_FILE_VERSION_MIN && f.version <= RWKV_FILE_VERSION_MAX,
             RWKV_ERROR_FILE | RWKV_ERROR_FILE_VERSION, false, "Unsupported file version %u", f.version);
    RB_CHECK(f.data_type < T_COUNT, RWKV_ERROR_FILE | RWKV_ERROR_DATA_TYPE, false,
             "Model data type out of range (%u > %u)", f.data_type, (unsigned) T_COUNT - 1);
    RB_CHECK(type_supported(f.data_type), RWKV_ERROR_FILE | RWKV_ERROR_DATA_TYPE, false,
             "Models in %s format cannot be loaded anymore because the format was removed", type_name(f.data_type));
    RB_CHECK(!type_quantized(f.data_type) || f.version == RWKV_FILE_VERSION_1, RWKV_ERROR_FILE | RWKV_ERROR_DATA_TYPE, false,
             "The quantized model file in %s format was created with an old version and can not be loaded", type_name(f.data_type));

    while (c.left > 0) {
        FileTensor t;
        uint32_t key_len = 0, ne0 = 0;
        RB_CHECK(c.u32(t.n_dims) && c.u32(key_len) && c.u32(t.type), RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_FILE_READ, false,
                 "Failed to read a model parameter");
        RB_CHECK(t.n_dims >= 1 && t.n_dims <= 3, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, false,
                 "Tensor has an invalid shape (%u dimensions)", t.n_dims);
        RB_CHECK(t.type < T_COUNT, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_DATA_TYPE, false, "Tensor data type out of range (%u)", t.type);
        RB_CHECK(type_supported(t.type), RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_DATA_TYPE, false,
                 "Tensor data type (%s) is no longer supported", type_name(t.type));
        for (uint32_t d = 0; d < t.n_dims; d++) {
            RB_CHECK(c.u32(ne0), RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_FILE_READ, false, "Failed to read a model parameter");
            t.ne[d] = ne0;
        }
        RB_CHECK(c.left >= key_len, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_KEY, false, "Failed to read a parameter name");
        t.name.assign((const char *) c.p, key_len);
        c.skip(key_len);
        // three unchecked 32-bit dimensions: their product must not wrap (a crafted header would otherwise
        // pass the size check below with a small nbytes while the loader copies ne[1]*ne[2] rows)
        {
            const unsigned __int128 n128 = (unsigned __int128) t.ne[0] * t.ne[1] * t.ne[2];
            RB_CHECK(t.ne[0] > 0 && t.ne[1] > 0 && t.ne[2] > 0 && n128 <= (unsigned __int128) f.map_size * 32, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, false,
                     "Tensor %s: dimensions (%llu, %llu, %llu) do not fit the file", t.name.c_str(), (unsigned long long) t.ne[0],
                     (unsigned long long) t.ne[1], (unsigned long long) t.ne[2]);
        }
        // rwkv_utilities.inc:1-4: nbytes = type_size * n / block
        RB_CHECK(t.ne[0] % type_block(t.type) == 0, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_SHAPE, false,
                 "Tensor %s: row length %llu is not a multiple of the block size", t.name.c_str(), (unsigned long long) t.ne[0]);
        t.nbytes = (size_t) (t.n_elements() / type_block(t.type)) * type_block_bytes(t.type);
        RB_CHECK(c.left >= t.nbytes, RWKV_ERROR_MODEL_PARAMS | RWKV_ERROR_DATA, false,
                 "Tensor %s: data runs past the end of the file", t.name.c_str());
        t.data = c.p;
        c.skip(t.nbytes);
        f.index[t.name] = f.tensors.size();
        f.tensors.push_back(std::move(t));
    }
    return true;
}

} // namespace rb
