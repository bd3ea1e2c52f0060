// Host-side CDF-5 ("64BIT_DATA") reader behind the ddpm_cdf5_* C ABI.
// Grammar: see pytorch_ddp_mnist_b200/cdf5.py (the same subset, written independently in C++ so
// the loader can pread() slabs straight into pinned memory without going through Python).
#include <errno.h>
#include <fcntl.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "../../include/ddpmnist_b200.h"

namespace ddpm { int set_error(const char* fmt, ...); }
using ddpm::set_error;

struct Var {
    std::string name;
    std::vector<int64_t> shape;
    int nc_type = 0;
    int elem = 0;
    int64_t vsize = 0, begin = 0;
};

struct ddpm_cdf5 {
    int fd = -1;
    int64_t file_size = 0;
    std::vector<std::pair<std::string, int64_t>> dims;
    std::vector<Var> vars;
};

namespace {

struct Cur {
    const uint8_t* p; size_t n, pos = 0; bool bad = false;
    bool need(size_t k) { if (pos + k > n) { bad = true; return false; } return true; }
    int32_t i32() { if (!need(4)) return 0; uint32_t v = 0; for (int i = 0; i < 4; ++i) v = (v << 8) | p[pos + i]; pos += 4; return (int32_t)v; }
    int64_t i64() { if (!need(8)) return 0; uint64_t v = 0; for (int i = 0; i < 8; ++i) v = (v << 8) | p[pos + i]; pos += 8; return (int64_t)v; }
    std::string name() {
        int64_t len = i64();
        if (bad || len < 0 || !need((size_t)((len + 3) & ~3LL))) { bad = true; return ""; }
        std::string s((const char*)p + pos, (size_t)len);
        pos += (size_t)((len + 3) & ~3LL);
        return s;
    }
};

int elem_size(int t) {
    switch (t) { case 1: case 2: case 7: return 1; case 3: case 8: return 2; case 4: case 5: case 9: return 4;
                 case 6: case 10: case 11: return 8; default: return 0; }
}

bool skip_atts(Cur& c) {
    int32_t tag = c.i32(); int64_t n = c.i64();
    if (c.bad) return false;
    if (tag == 0) return n == 0;
    if (tag != 0x0C) return false;
    for (int64_t i = 0; i < n; ++i) {
        c.name();
        int32_t t = c.i32(); int64_t ne = c.i64();
        int es = elem_size(t);
        if (c.bad || es == 0 || ne < 0) return false;
        size_t bytes = (size_t)((ne * es + 3) & ~3LL);
        if (!c.need(bytes)) return false;
        c.pos += bytes;
    }
    return true;
}

// returns 0 ok, 1 malformed, 2 need more bytes
int parse(ddpm_cdf5* f, const uint8_t* buf, size_t n) {
    if (n < 4 || memcmp(buf, "CDF", 3) != 0) { set_error("not a netCDF classic file (bad magic)"); return 1; }
    if (buf[3] != 5) { set_error("unsupported netCDF classic version %d (only CDF-5 / 64BIT_DATA)", (int)buf[3]); return 1; }
    Cur c{buf, n}; c.pos = 4;
    int64_t numrecs = c.i64(); (void)numrecs;
    int32_t tag = c.i32(); int64_t cnt = c.i64();
    if (c.bad) return 2;
    f->dims.clear(); f->vars.clear();
    if (tag == 0x0A) {
        for (int64_t i = 0; i < cnt; ++i) { std::string nm = c.name(); int64_t len = c.i64(); if (c.bad) return 2; f->dims.push_back({nm, len}); }
    } else if (tag != 0 || cnt != 0) { set_error("CDF-5: bad dim_list tag %#x", tag); return 1; }
    if (!skip_atts(c)) { if (c.bad) return 2; set_error("CDF-5: bad global attribute list"); return 1; }
    tag = c.i32(); cnt = c.i64();
    if (c.bad) return 2;
    if (tag == 0x0B) {
        for (int64_t i = 0; i < cnt; ++i) {
            Var v; v.name = c.name();
            int64_t nd = c.i64();
            if (c.bad) return 2;
            if (nd < 0 || nd > 8) { set_error("CDF-5: variable %s has %lld dimensions", v.name.c_str(), (long long)nd); return 1; }
            for (int64_t d = 0; d < nd; ++d) {
                int64_t id = c.i64();
                if (c.bad) return 2;
                if (id < 0 || id >= (int64_t)f->dims.size()) { set_error("CDF-5: variable %s references dimension id %lld", v.name.c_str(), (long long)id); return 1; }
                v.shape.push_back(f->dims[(size_t)id].second);
            }
            if (!skip_atts(c)) { if (c.bad) return 2; set_error("CDF-5: bad attribute list of %s", v.name.c_str()); return 1; }
            v.nc_type = c.i32(); v.vsize = c.i64(); v.begin = c.i64();
            if (c.bad) return 2;
            v.elem = elem_size(v.nc_type);
            if (v.elem == 0) { set_error("CDF-5: unknown nc_type %d for %s", v.nc_type, v.name.c_str()); return 1; }
            f->vars.push_back(v);
        }
    } else if (tag != 0 || cnt != 0) { set_error("CDF-5: bad var_list tag %#x", tag); return 1; }
    return 0;
}

const Var* find(const ddpm_cdf5* f, const char* name) {
    for (const Var& v : f->vars) if (v.name == name) return &v;
    return nullptr;
}

}  // namespace

extern "C" {

int ddpm_cdf5_open(const char* path, ddpm_cdf5** out) {
    if (!path || !out) return set_error("ddpm_cdf5_open: null argument");
    *out = nullptr;
    int fd = open(path, O_RDONLY);
    if (fd < 0) return set_error("ddpm_cdf5_open: cannot open %s: %s", path, strerror(errno));
    struct stat st;
    if (fstat(fd, &st) != 0) { close(fd); return set_error("ddpm_cdf5_open: fstat failed: %s", strerror(errno)); }
    ddpm_cdf5* f = new ddpm_cdf5();
    f->fd = fd; f->file_size = st.st_size;
    size_t want = 8192;
    for (;;) {
        size_t n = want < (size_t)st.st_size ? want : (size_t)st.st_size;
        std::vector<uint8_t> buf(n);
        ssize_t got = pread(fd, buf.data(), n, 0);
        if (got < 0 || (size_t)got != n) { close(fd); delete f; return set_error("ddpm_cdf5_open: short read of header"); }
        int rc = parse(f, buf.data(), n);
        if (rc == 0) break;
        if (rc == 1) { close(fd); delete f; return 1; }
        if (n == (size_t)st.st_size) { close(fd); delete f; return set_error("CDF-5: truncated header"); }
        want *= 4;
    }
    for (const Var& v : f->vars) {
        int64_t bytes = v.elem;
        for (int64_t s : v.shape) bytes *= s;
        if (v.begin < 0 || v.begin + bytes > f->file_size) {
            close(fd); delete f;
            return set_error("CDF-5: variable %s [%lld, +%lld) lies outside the file (%lld bytes)", v.name.c_str(),
                             (long long)v.begin, (long long)bytes, (long long)st.st_size);
        }
    }
    *out = f;
    return 0;
}

void ddpm_cdf5_close(ddpm_cdf5* f) {
    if (!f) return;
    if (f->fd >= 0) close(f->fd);
    delete f;
}

int ddpm_cdf5_inq_var(const ddpm_cdf5* f, const char* name, int* ndims, int64_t* dims,
                      int* nc_type, int* elem_size_out, int64_t* begin) {
    if (!f || !name) return set_error("ddpm_cdf5_inq_var: null argument");
    const Var* v = find(f, name);
    if (!v) return set_error("ddpm_cdf5_inq_var: no variable named '%s'", name);
    if (ndims) *ndims = (int)v->shape.size();
    if (dims) for (size_t i = 0; i < v->shape.size(); ++i) dims[i] = v->shape[i];
    if (nc_type) *nc_type = v->nc_type;
    if (elem_size_out) *elem_size_out = v->elem;
    if (begin) *begin = v->begin;
    return 0;
}

int ddpm_cdf5_read_rows(const ddpm_cdf5* f, const char* name, int64_t row0, int64_t nrows, void* h_dst) {
    if (!f || !name) return set_error("ddpm_cdf5_read_rows: null argument");
    const Var* v = find(f, name);
    if (!v) return set_error("ddpm_cdf5_read_rows: no variable named '%s'", name);
    if (v->shape.empty()) return set_error("ddpm_cdf5_read_rows: '%s' is a scalar", name);
    if (row0 < 0 || nrows < 0 || row0 + nrows > v->shape[0])
        return set_error("ddpm_cdf5_read_rows: rows [%lld, %lld) out of range for '%s' (%lld rows)",
                         (long long)row0, (long long)(row0 + nrows), name, (long long)v->shape[0]);
    if (nrows == 0) return 0;
    if (!h_dst) return set_error("ddpm_cdf5_read_rows: null destination");
    int64_t row_bytes = v->elem;
    for (size_t i = 1; i < v->shape.size(); ++i) row_bytes *= v->shape[i];
    int64_t off = v->begin + row0 * row_bytes, left = nrows * row_bytes;
    char* dst = (char*)h_dst;
    while (left > 0) {
        size_t chunk = left > (int64_t)(1 << 30) ? (size_t)(1 << 30) : (size_t)left;
        ssize_t got = pread(f->fd, dst, chunk, off);
        if (got <= 0) return set_error("ddpm_cdf5_read_rows: pread failed at offset %lld: %s", (long long)off,
                                       got < 0 ? strerror(errno) : "unexpected end of file");
        dst += got; off += got; left -= got;
    }
    return 0;
}

}  // extern "C"
