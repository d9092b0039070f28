// Native .npz model loader: zip central directory + stored / deflate members + .npy headers, then the
// array conventions of the reference's loader so that a licensed SMPL model.npz gives the same model as ark::AvatarModel:
//   external/avatar/cnpy.cpp:67-301        npy header / npz member parsing (little-endian, word size, fortran_order, shape)
//   external/avatar/Util.cpp:249-298       loadFloatMatrix / loadUintMatrix: element width 4 or 8, values cast to double / uint32,
//                                          fortran_order honoured for the (rows x cols) view the CALLER asks for
//   external/avatar/AvatarModel.cpp:23-102 which arrays, which views: kintree_table (2 x J, row 0), v_template read FLAT as
//                                          1 x 3V (so fortran_order is ignored there), f (F x 3), J_regressor (J x V), weights (V x J),
//                                          shapedirs as a (3V x S) matrix
// Host-only code (no kernels); lives in the CUDA library because b3d_model_load_npz ends in b3d_model_create.
#include <zlib.h>

#include <algorithm>
#include <cstdlib>
#include <map>

#include "internal.hpp"

namespace b3d {
namespace {

struct NpyArray {
    std::vector<unsigned char> data;     // raw little-endian elements
    std::vector<int64_t> shape;
    int word = 0;                        // bytes per element
    char kind = 0;                       // 'f', 'i', 'u', 'b', ...
    bool fortran = false;
    size_t count() const { size_t n = 1; for (int64_t d : shape) n *= (size_t)d; return n; }
};

uint16_t rd16(const unsigned char* p) { return (uint16_t)(p[0] | (p[1] << 8)); }
uint32_t rd32(const unsigned char* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
uint64_t rd64(const unsigned char* p) { return (uint64_t)rd32(p) | ((uint64_t)rd32(p + 4) << 32); }

struct ZipEntry { uint64_t csize, usize, offset; int method; };

// central directory -> name -> entry
int zip_index(FILE* f, std::map<std::string, ZipEntry>& out) {
    if (fseek(f, 0, SEEK_END)) return B3D_EFORMAT;
    const long fsize = ftell(f);
    if (fsize < 22) return B3D_EFORMAT;
    const long tail = fsize < 65557 ? fsize : 65557;
    std::vector<unsigned char> buf((size_t)tail);
    fseek(f, fsize - tail, SEEK_SET);
    if (fread(buf.data(), 1, (size_t)tail, f) != (size_t)tail) return B3D_EFORMAT;
    long eocd = -1;
    for (long i = tail - 22; i >= 0; --i)
        if (rd32(&buf[(size_t)i]) == 0x06054b50u) { eocd = i; break; }
    if (eocd < 0) return B3D_EFORMAT;
    uint64_t n = rd16(&buf[(size_t)eocd + 10]), cd_size = rd32(&buf[(size_t)eocd + 12]), cd_off = rd32(&buf[(size_t)eocd + 16]);
    if (n == 0xFFFFu || cd_size == 0xFFFFFFFFu || cd_off == 0xFFFFFFFFu) {     // ZIP64: locator sits right before the EOCD
        if (eocd < 20 || rd32(&buf[(size_t)eocd - 20]) != 0x07064b50u) return B3D_EFORMAT;
        const uint64_t e64 = rd64(&buf[(size_t)eocd - 20 + 8]);
        unsigned char h[56];
        if (fseek(f, (long)e64, SEEK_SET) || fread(h, 1, 56, f) != 56 || rd32(h) != 0x06064b50u) return B3D_EFORMAT;
        n = rd64(h + 32); cd_size = rd64(h + 40); cd_off = rd64(h + 48);
    }
    if (cd_off + cd_size > (uint64_t)fsize) return B3D_EFORMAT;
    std::vector<unsigned char> cd((size_t)cd_size);
    if (fseek(f, (long)cd_off, SEEK_SET) || fread(cd.data(), 1, cd.size(), f) != cd.size()) return B3D_EFORMAT;
    size_t p = 0;
    for (uint64_t i = 0; i < n; ++i) {
        if (p + 46 > cd.size() || rd32(&cd[p]) != 0x02014b50u) return B3D_EFORMAT;
        ZipEntry e;
        e.method = rd16(&cd[p + 10]);
        e.csize = rd32(&cd[p + 20]); e.usize = rd32(&cd[p + 24]); e.offset = rd32(&cd[p + 42]);
        const size_t nl = rd16(&cd[p + 28]), xl = rd16(&cd[p + 30]), cl = rd16(&cd[p + 32]);
        if (p + 46 + nl + xl + cl > cd.size()) return B3D_EFORMAT;
        const std::string name((const char*)&cd[p + 46], nl);
        size_t x = p + 46 + nl;
        const size_t xend = x + xl;
        while (x + 4 <= xend) {                                  // zip64 extended information: only the saturated fields, in order
            const uint16_t id = rd16(&cd[x]), len = rd16(&cd[x + 2]);
            if (id == 0x0001) {
                size_t q = x + 4;
                if (e.usize == 0xFFFFFFFFu && q + 8 <= xend) { e.usize = rd64(&cd[q]); q += 8; }
                if (e.csize == 0xFFFFFFFFu && q + 8 <= xend) { e.csize = rd64(&cd[q]); q += 8; }
                if (e.offset == 0xFFFFFFFFu && q + 8 <= xend) { e.offset = rd64(&cd[q]); q += 8; }
            }
            x += 4 + (size_t)len;
        }
        out[name] = e;
        p += 46 + nl + xl + cl;
    }
    return B3D_OK;
}

int zip_member(FILE* f, const ZipEntry& e, std::vector<unsigned char>& out) {
    unsigned char h[30];
    if (fseek(f, (long)e.offset, SEEK_SET) || fread(h, 1, 30, f) != 30 || rd32(h) != 0x04034b50u) return B3D_EFORMAT;
    if (fseek(f, (long)(e.offset + 30 + rd16(h + 26) + rd16(h + 28)), SEEK_SET)) return B3D_EFORMAT;
    {   // sizes come from the file: bound them by what the file can hold (deflate expands at most ~1032 : 1)
        const long here = ftell(f);
        if (here < 0 || fseek(f, 0, SEEK_END)) return B3D_EFORMAT;
        const long fsize = ftell(f);
        if (fseek(f, here, SEEK_SET)) return B3D_EFORMAT;
        if (e.csize > (uint64_t)(fsize - here) || e.usize > 1040ull * (e.csize + 64)) return B3D_EFORMAT;
    }
    std::vector<unsigned char> comp((size_t)e.csize);
    if (fread(comp.data(), 1, comp.size(), f) != comp.size()) return B3D_EFORMAT;
    if (e.method == 0) {
        if (e.csize != e.usize) return B3D_EFORMAT;
        out.swap(comp);
        return B3D_OK;
    }
    if (e.method != 8) return B3D_EFORMAT;
    out.resize((size_t)e.usize);
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (inflateInit2(&zs, -MAX_WBITS) != Z_OK) return B3D_ENOMEM;
    int rc = Z_OK;
    size_t in_done = 0, out_done = 0;
    while (rc == Z_OK) {                                         // avail_in / avail_out are 32-bit: feed in chunks
        const size_t in_chunk = std::min<size_t>(comp.size() - in_done, 1u << 30), out_chunk = std::min<size_t>(out.size() - out_done, 1u << 30);
        zs.next_in = comp.data() + in_done; zs.avail_in = (uInt)in_chunk;
        zs.next_out = out.data() + out_done; zs.avail_out = (uInt)out_chunk;
        rc = inflate(&zs, Z_NO_FLUSH);
        in_done += in_chunk - zs.avail_in;
        out_done += out_chunk - zs.avail_out;
        if (rc == Z_OK && in_chunk - zs.avail_in == 0 && out_chunk - zs.avail_out == 0) break;   // no progress
    }
    inflateEnd(&zs);
    return (rc == Z_STREAM_END && out_done == out.size()) ? B3D_OK : B3D_EFORMAT;
}

// .npy: "\x93NUMPY" major minor hlen header-dict data (cnpy.cpp:67-118)
int parse_npy(const std::vector<unsigned char>& raw, NpyArray& a) {
    if (raw.size() < 10 || memcmp(raw.data(), "\x93NUMPY", 6) != 0) return B3D_EFORMAT;
    const int major = raw[6];
    size_t hlen, hoff;
    if (major == 1) { hlen = rd16(&raw[8]); hoff = 10; }
    else if (major == 2 || major == 3) { if (raw.size() < 12) return B3D_EFORMAT; hlen = rd32(&raw[8]); hoff = 12; }
    else return B3D_EFORMAT;
    if (hoff + hlen > raw.size()) return B3D_EFORMAT;
    const std::string h((const char*)&raw[hoff], hlen);
    size_t p = h.find("'descr'");
    if (p == std::string::npos) return B3D_EFORMAT;
    p = h.find('\'', h.find(':', p));
    if (p == std::string::npos || p + 3 >= h.size()) return B3D_EFORMAT;
    const char endian = h[p + 1];
    if (endian != '<' && endian != '|' && endian != '=') return B3D_EFORMAT;        // big-endian files are not supported (cnpy asserts too)
    a.kind = h[p + 2];
    a.word = atoi(h.c_str() + p + 3);
    p = h.find("'fortran_order'");
    if (p == std::string::npos) return B3D_EFORMAT;
    {
        const size_t v = h.find_first_not_of(" :", p + 15);
        if (v == std::string::npos) return B3D_EFORMAT;
        a.fortran = h.compare(v, 4, "True") == 0;
    }
    p = h.find("'shape'");
    if (p == std::string::npos) return B3D_EFORMAT;
    const size_t lo = h.find('(', p), hi = h.find(')', p);
    if (lo == std::string::npos || hi == std::string::npos || hi < lo) return B3D_EFORMAT;
    a.shape.clear();
    for (size_t q = lo + 1; q < hi;) {
        while (q < hi && (h[q] == ' ' || h[q] == ',')) ++q;
        if (q >= hi) break;
        char* end = nullptr;
        const long long dim = strtoll(h.c_str() + q, &end, 10);
        if (end == h.c_str() + q || dim < 0) return B3D_EFORMAT;      // not a number / negative dimension
        a.shape.push_back(dim);
        q = (size_t)(end - h.c_str());
    }
    {   // element count must not overflow and must fit the payload
        unsigned long long cnt = 1;
        for (long long d : a.shape) {
            if (d != 0 && cnt > (1ull << 40) / (unsigned long long)d) return B3D_EFORMAT;
            cnt *= (unsigned long long)d;
        }
    }
    if (a.word <= 0 || raw.size() - hoff - hlen < a.count() * (size_t)a.word) return B3D_EFORMAT;
    a.data.assign(raw.begin() + (long)(hoff + hlen), raw.begin() + (long)(hoff + hlen + a.count() * (size_t)a.word));
    return B3D_OK;
}

int npz_get(const char* path, const char* name, NpyArray& a) {
    FILE* f = fopen(path, "rb");
    if (!f) return set_error(B3D_ENOFILE, "cannot open %s", path);
    std::map<std::string, ZipEntry> idx;
    int rc = zip_index(f, idx);
    if (rc == B3D_OK) {
        auto it = idx.find(std::string(name) + ".npy");
        if (it == idx.end()) rc = B3D_ENOFILE;
        else {
            try {       // nothing may escape through the extern "C" callers
                std::vector<unsigned char> raw;
                rc = zip_member(f, it->second, raw);
                if (rc == B3D_OK) rc = parse_npy(raw, a);
            } catch (const std::bad_alloc&) {
                rc = B3D_ENOMEM;
            } catch (const std::exception&) {
                rc = B3D_EFORMAT;
            }
        }
    }
    fclose(f);
    if (rc == B3D_ENOFILE) return set_error(rc, "%s: no array '%s'", path, name);
    if (rc == B3D_ENOMEM) return set_error(rc, "%s: out of memory reading '%s'", path, name);
    if (rc != B3D_OK) return set_error(rc, "%s: malformed zip / npy data for '%s'", path, name);
    return B3D_OK;
}

// Util.cpp:249-271: (rows x cols) view, row-major unless fortran_order, float32 / float64 -> double
int float_matrix(const NpyArray& a, size_t r, size_t c, std::vector<double>& out, const char* what) {
    if (a.kind != 'f' || (a.word != 4 && a.word != 8) || a.count() != r * c) return set_error(B3D_EFORMAT, "%s: expected %zu float32/float64 values", what, r * c);
    out.resize(r * c);
    for (size_t i = 0; i < r; ++i)
        for (size_t j = 0; j < c; ++j) {
            const size_t src = a.fortran ? i + r * j : i * c + j;
            double v;
            if (a.word == 4) { float t; memcpy(&t, &a.data[src * 4], 4); v = (double)t; }
            else memcpy(&v, &a.data[src * 8], 8);
            out[i * c + j] = v;
        }
    return B3D_OK;
}

// Util.cpp:272-298: 32- or 64-bit integers -> uint32 (a 64-bit -1 becomes 0xFFFFFFFF like Eigen's cast)
int uint_matrix(const NpyArray& a, size_t r, size_t c, std::vector<uint32_t>& out, const char* what) {
    if ((a.kind != 'i' && a.kind != 'u') || (a.word != 4 && a.word != 8) || a.count() != r * c)
        return set_error(B3D_EFORMAT, "%s: expected %zu 32/64-bit integers", what, r * c);
    out.resize(r * c);
    for (size_t i = 0; i < r; ++i)
        for (size_t j = 0; j < c; ++j) {
            const size_t src = a.fortran ? i + r * j : i * c + j;
            out[i * c + j] = a.word == 4 ? rd32(&a.data[src * 4]) : (uint32_t)rd64(&a.data[src * 8]);
        }
    return B3D_OK;
}

bool shape_is(const NpyArray& a, std::initializer_list<int64_t> s) {      // Util.cpp:300-308 assertShape (-1 = any)
    if (a.shape.size() != s.size()) return false;
    size_t k = 0;
    for (int64_t d : s) { if (d >= 0 && a.shape[k] != d) return false; ++k; }
    return true;
}

}  // namespace
}  // namespace b3d

using namespace b3d;

extern "C" {

int b3d_npz_shape(const char* path, const char* name, int32_t* ndim, int64_t* shape4, int32_t* word_size, int32_t* fortran_order) {
    if (!path || !name) return set_error(B3D_EINVAL, "b3d_npz_shape: null argument");
    NpyArray a;
    if (int rc = npz_get(path, name, a)) return rc;
    if (a.shape.size() > 4) return set_error(B3D_EFORMAT, "%s: '%s' has more than 4 dimensions", path, name);
    if (ndim) *ndim = (int32_t)a.shape.size();
    if (shape4) for (size_t i = 0; i < a.shape.size(); ++i) shape4[i] = a.shape[i];
    if (word_size) *word_size = a.word;
    if (fortran_order) *fortran_order = a.fortran ? 1 : 0;
    return B3D_OK;
}

int b3d_npz_float_matrix(const char* path, const char* name, int64_t rows, int64_t cols, double* out) {
    if (!path || !name || !out || rows <= 0 || cols <= 0) return set_error(B3D_EINVAL, "b3d_npz_float_matrix: bad argument");
    NpyArray a;
    if (int rc = npz_get(path, name, a)) return rc;
    std::vector<double> m;
    if (int rc = float_matrix(a, (size_t)rows, (size_t)cols, m, name)) return rc;
    memcpy(out, m.data(), sizeof(double) * m.size());
    return B3D_OK;
}

int b3d_npz_uint_matrix(const char* path, const char* name, int64_t rows, int64_t cols, uint32_t* out) {
    if (!path || !name || !out || rows <= 0 || cols <= 0) return set_error(B3D_EINVAL, "b3d_npz_uint_matrix: bad argument");
    NpyArray a;
    if (int rc = npz_get(path, name, a)) return rc;
    std::vector<uint32_t> m;
    if (int rc = uint_matrix(a, (size_t)rows, (size_t)cols, m, name)) return rc;
    memcpy(out, m.data(), sizeof(uint32_t) * m.size());
    return B3D_OK;
}

int b3d_model_load_npz(const char* path, uint32_t flags, b3d_model** out) {
    if (!path || !out) return set_error(B3D_EINVAL, "b3d_model_load_npz: null argument");
    NpyArray vt, kt, fa, jr, wt, sd, pd;
    if (int rc = npz_get(path, "v_template", vt)) return rc;
    if (int rc = npz_get(path, "kintree_table", kt)) return rc;
    if (int rc = npz_get(path, "f", fa)) return rc;
    if (int rc = npz_get(path, "J_regressor", jr)) return rc;
    if (int rc = npz_get(path, "weights", wt)) return rc;
    if (int rc = npz_get(path, "shapedirs", sd)) return rc;
    if (vt.shape.size() < 1 || kt.shape.size() != 2 || fa.shape.size() < 1 || sd.shape.size() != 3)
        return set_error(B3D_EFORMAT, "%s: unexpected array ranks", path);
    const int64_t V = vt.shape[0], J = kt.shape[1], F = fa.shape[0], S = sd.shape[2];       // AvatarModel.cpp:26-29
    if (!shape_is(vt, {V, 3}) || !shape_is(kt, {2, J}) || !shape_is(fa, {F, 3}) || !shape_is(jr, {J, V}) || !shape_is(wt, {V, J}) ||
        !shape_is(sd, {V, 3, S}))
        return set_error(B3D_EFORMAT, "%s: array shapes do not fit together (V %lld, J %lld, F %lld, S %lld)", path, (long long)V, (long long)J,
                         (long long)F, (long long)S);
    std::vector<uint32_t> kin, faces;
    std::vector<double> v, jreg, w, shp, pose;
    if (int rc = uint_matrix(kt, 2, (size_t)J, kin, "kintree_table")) return rc;
    if (int rc = float_matrix(vt, 1, (size_t)V * 3, v, "v_template")) return rc;            // read flat (:45-48)
    if (int rc = uint_matrix(fa, (size_t)F, 3, faces, "f")) return rc;
    if (int rc = float_matrix(jr, (size_t)J, (size_t)V, jreg, "J_regressor")) return rc;
    if (int rc = float_matrix(wt, (size_t)V, (size_t)J, w, "weights")) return rc;
    if (int rc = float_matrix(sd, (size_t)V * 3, (size_t)S, shp, "shapedirs")) return rc;   // (3V x S) view (:98-99)
    bool have_pd = false;
    if (flags & B3D_NPZ_WITH_POSEDIRS) {
        if (int rc = npz_get(path, "posedirs", pd)) return rc;
        if (!shape_is(pd, {V, 3, 9 * (J - 1)})) return set_error(B3D_EFORMAT, "%s: posedirs is not V x 3 x %lld", path, (long long)(9 * (J - 1)));
        if (int rc = float_matrix(pd, (size_t)V * 3, (size_t)(9 * (J - 1)), pose, "posedirs")) return rc;
        have_pd = true;
    }
    b3d_model_desc d;
    memset(&d, 0, sizeof(d));
    d.num_verts = (int32_t)V; d.num_joints = (int32_t)J; d.num_shape = (int32_t)S; d.num_faces = (int32_t)F;
    d.dtype = B3D_F64;
    d.v_template = v.data(); d.shapedirs = shp.data(); d.J_regressor = jreg.data(); d.weights = w.data();
    d.kintree_row0 = kin.data();                 // row 0 of the 2 x J table = parents
    d.faces = faces.data();
    d.posedirs = have_pd ? pose.data() : nullptr;
    return b3d_model_create(&d, out);
}

}  // extern "C"
