// uniffi_abi.cpp -- libslimg_ffi's UniFFI scaffolding symbols over the C ABI of this library (include/slimg_uniffi.h).
//
// What the reference generates from crates/slimg-ffi/src/lib.rs with `uniffi::setup_scaffolding!()`; the wire format
// is the one the checked-in Kotlin binding lifts and lowers (bindings/kotlin/.../slimg_ffi.kt).  Host code only:
// the pixel work goes through slimg_b200_{resize,crop,extend,pipeline}, i.e. the CUDA kernels.
#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/slimg_uniffi.h"
#include "host.h"

namespace {

// ---- big-endian reader / writer over a byte buffer (slimg_ffi.kt:82-85: ByteOrder.BIG_ENDIAN) ----------------
struct Reader {
    const uint8_t *p;
    size_t n, i = 0;
    bool ok = true;
    Reader(const RustBuffer &b) : p(b.data), n(size_t(b.len)) {}
    bool need(size_t k) {
        if (!ok || n - i < k) { ok = false; return false; }
        return true;
    }
    uint8_t u8() { return need(1) ? p[i++] : 0; }
    uint32_t u32() {
        if (!need(4)) return 0;
        const uint32_t v = (uint32_t(p[i]) << 24) | (uint32_t(p[i + 1]) << 16) | (uint32_t(p[i + 2]) << 8) | p[i + 3];
        i += 4;
        return v;
    }
    int32_t i32() { return int32_t(u32()); }
    double f64() {
        if (!need(8)) return 0.0;
        uint64_t v = 0;
        for (int k = 0; k < 8; k++) v = (v << 8) | p[i + k];
        i += 8;
        double d;
        std::memcpy(&d, &v, 8);
        return d;
    }
    const uint8_t *bytes(size_t k) {
        if (!need(k)) return nullptr;
        const uint8_t *q = p + i;
        i += k;
        return q;
    }
    bool done() const { return ok && i == n; }
};

struct Writer {
    std::vector<uint8_t> b;
    void u8(uint8_t v) { b.push_back(v); }
    void u32(uint32_t v) { for (int s = 24; s >= 0; s -= 8) b.push_back(uint8_t(v >> s)); }
    void i32(int32_t v) { u32(uint32_t(v)); }
    void bytes(const void *p, size_t n) { const uint8_t *q = static_cast<const uint8_t *>(p); b.insert(b.end(), q, q + n); }
    void str(const std::string &s) { i32(int32_t(s.size())); bytes(s.data(), s.size()); }
};

RustBuffer make_buffer(const void *p, size_t n) {
    RustBuffer r;
    r.capacity = n; r.len = n;
    r.data = n ? static_cast<uint8_t *>(std::malloc(n)) : nullptr;
    if (n && !r.data) { r.capacity = r.len = 0; return r; }
    if (n && p) std::memcpy(r.data, p, n);
    return r;
}
RustBuffer take(const Writer &w) { return make_buffer(w.b.data(), w.b.size()); }
const RustBuffer kEmpty = {0, 0, nullptr};
void release(RustBuffer &b) { std::free(b.data); b = kEmpty; }

void set_ok(RustCallStatus *st) { if (st) st->code = 0; }

// SlimgError: i32 tag (the slimg_status value) + one string field (slimg_ffi.kt:1961-1997)
RustBuffer raise(RustCallStatus *st, int32_t tag, const std::string &msg) {
    if (st) {
        Writer w;
        w.i32(tag);
        w.str(msg);
        st->code = 1;
        st->error_buf = take(w);
    }
    return kEmpty;
}
RustBuffer raise_last(RustCallStatus *st, slimg_status s) { return raise(st, int32_t(s), slimg_b200_last_error()); }
// a buffer that does not parse is a bug of the caller's binding: UniFFI reports those as "unexpected" (code 2, a string)
RustBuffer panic(RustCallStatus *st, const char *what) {
    if (st) { st->code = 2; st->error_buf = make_buffer(what, std::strlen(what)); }
    return kEmpty;
}

// ---- lifting ------------------------------------------------------------------------------------------------
bool read_image(Reader &r, slimg_image &img) {
    img.width = r.u32(); img.height = r.u32();
    const int32_t len = r.i32();
    if (!r.ok || len < 0) return false;
    img.data = const_cast<uint8_t *>(r.bytes(size_t(len)));
    img.len = size_t(len);
    return r.ok;
}
bool read_resize(Reader &r, slimg_resize_mode &m) {
    m = slimg_resize_mode{};
    m.tag = r.i32();
    switch (m.tag) {
        case SLIMG_RESIZE_WIDTH: case SLIMG_RESIZE_HEIGHT: m.a = r.u32(); break;
        case SLIMG_RESIZE_EXACT: case SLIMG_RESIZE_FIT: m.a = r.u32(); m.b = r.u32(); break;
        case SLIMG_RESIZE_SCALE: m.factor = r.f64(); break;
        default: return false;
    }
    return r.ok;
}
bool read_crop(Reader &r, slimg_crop_mode &m) {
    m = slimg_crop_mode{};
    m.tag = r.i32();
    if (m.tag == SLIMG_CROP_REGION) { m.x = r.u32(); m.y = r.u32(); m.width = r.u32(); m.height = r.u32(); }
    else if (m.tag == SLIMG_CROP_ASPECT_RATIO) { m.width = r.u32(); m.height = r.u32(); }
    else return false;
    return r.ok;
}
bool read_extend(Reader &r, slimg_extend_mode &m) {
    m = slimg_extend_mode{};
    m.tag = r.i32();
    if (m.tag != SLIMG_EXTEND_ASPECT_RATIO && m.tag != SLIMG_EXTEND_SIZE) return false;
    m.width = r.u32(); m.height = r.u32();
    return r.ok;
}
bool read_fill(Reader &r, slimg_fill_color &f) {
    f = slimg_fill_color{};
    f.tag = r.i32();
    if (f.tag == SLIMG_FILL_SOLID) { f.r = r.u8(); f.g = r.u8(); f.b = r.u8(); f.a = r.u8(); }
    else if (f.tag != SLIMG_FILL_TRANSPARENT) return false;
    return r.ok;
}
bool read_format(Reader &r, int32_t &f) { f = r.i32(); return r.ok && f >= 1 && f <= 6; }
bool read_option_tag(Reader &r, bool &present) {
    const uint8_t t = r.u8();
    present = t == 1;
    return r.ok && t <= 1;
}

// ---- lowering -----------------------------------------------------------------------------------------------
RustBuffer lower_image(const slimg_image &img) {
    Writer w;
    w.b.reserve(12 + img.len);
    w.u32(img.width); w.u32(img.height); w.i32(int32_t(img.len));
    w.bytes(img.data, img.len);
    return take(w);
}

const char *kExt[7] = {"", "jpg", "png", "webp", "avif", "jxl", "qoi"};          // format.rs:58-67

std::string lower_ascii(std::string s) {
    for (char &c : s) if (c >= 'A' && c <= 'Z') c = char(c - 'A' + 'a');
    return s;
}
// std::path::Path::extension: the part of the FILE NAME after its last '.', none for ".hidden" or no dot
bool path_extension(const std::string &path, std::string &ext) {
    size_t end = path.size();
    while (end > 1 && path[end - 1] == '/') end--;
    const size_t slash = path.rfind('/', end ? end - 1 : 0);
    const size_t start = slash == std::string::npos ? 0 : slash + 1;
    const std::string name = path.substr(start, end - start);
    if (name.empty() || name == "..") return false;
    const size_t dot = name.rfind('.');
    if (dot == std::string::npos || dot == 0) return false;
    ext = name.substr(dot + 1);
    return true;
}
int32_t format_from_ext(const std::string &path) {                                 // format.rs:16-27
    std::string ext;
    if (!path_extension(path, ext)) return 0;
    ext = lower_ascii(ext);
    if (ext == "jpg" || ext == "jpeg") return 1;
    if (ext == "png") return 2;
    if (ext == "webp") return 3;
    if (ext == "avif") return 4;
    if (ext == "jxl") return 5;
    if (ext == "qoi") return 6;
    return 0;
}
int32_t format_from_magic(const uint8_t *d, size_t n) {                            // format.rs:30-55
    if (n >= 3 && d[0] == 0xFF && d[1] == 0xD8 && d[2] == 0xFF) return 1;
    if (n >= 4 && d[0] == 0x89 && d[1] == 0x50 && d[2] == 0x4E && d[3] == 0x47) return 2;
    if (n >= 12 && !std::memcmp(d, "RIFF", 4) && !std::memcmp(d + 8, "WEBP", 4)) return 3;
    if (n >= 12 && !std::memcmp(d + 4, "ftyp", 4) && (!std::memcmp(d + 8, "avif", 4) || !std::memcmp(d + 8, "avis", 4))) return 4;
    if (n >= 2 && d[0] == 0xFF && d[1] == 0x0A) return 5;
    if (n >= 8 && d[0] == 0 && d[1] == 0 && d[2] == 0 && d[3] == 0x0C && !std::memcmp(d + 4, "JXL ", 4)) return 5;
    if (n >= 4 && !std::memcmp(d, "qoif", 4)) return 6;
    return 0;
}
// Path::with_extension: replace (or add) the extension of the file name
std::string with_extension(const std::string &path, const char *ext) {
    std::string old;
    std::string base = path;
    if (path_extension(path, old)) base = path.substr(0, path.size() - old.size() - 1);
    return base + "." + ext;
}
std::string file_stem(const std::string &path) {
    const size_t slash = path.rfind('/');
    std::string name = slash == std::string::npos ? path : path.substr(slash + 1);
    const size_t dot = name.rfind('.');
    if (dot != std::string::npos && dot != 0) name = name.substr(0, dot);
    return name;
}

struct Hooks {
    slimg_encode_hook enc = nullptr;
    slimg_decode_hook dec = nullptr;
    void *user = nullptr;
    // Format::Jpeg when a coefficient-taking entropy coder is installed: convert never downloads pixels
    slimg_jpeg_coef_hook jpeg = nullptr;
    void *jpeg_user = nullptr;
    int32_t jpeg_tables = SLIMG_QT_MOZJPEG;
    uint32_t jpeg_flags = 0;
};
Hooks g_hooks;

RustBuffer lower_pipeline_result(const uint8_t *data, size_t len, int32_t format, uint32_t w, uint32_t h) {
    Writer o;
    o.b.reserve(16 + len);
    o.i32(int32_t(len)); o.bytes(data, len); o.i32(format); o.u32(w); o.u32(h);      // lib.rs:174-184
    return take(o);
}

RustBuffer encode_with_hook(RustCallStatus *st, const slimg_image &pixels, int32_t format, uint8_t quality) {
    if (!g_hooks.enc)
        return raise(st, SLIMG_ERR_ENCODE, std::string("no encoder linked for ") + kExt[format] +
                                               ": entropy coding stays on the reference codecs (slimg_b200_set_codec_hooks)");
    uint8_t *out = nullptr;
    size_t out_len = 0;
    const char *msg = nullptr;
    const int32_t rc = g_hooks.enc(g_hooks.user, &pixels, format, quality, &out, &out_len, &msg);
    if (rc != SLIMG_OK) { std::free(out); return raise(st, rc, msg ? msg : "encoder hook failed"); }
    RustBuffer r = lower_pipeline_result(out, out_len, format, pixels.width, pixels.height);
    std::free(out);
    set_ok(st);
    return r;
}

}  // namespace

extern "C" {

// ---- buffers ------------------------------------------------------------------------------------------------
RustBuffer ffi_slimg_ffi_rustbuffer_alloc(uint64_t size, RustCallStatus *status) {
    set_ok(status);
    RustBuffer r = make_buffer(nullptr, size_t(size));
    if (size && !r.data) return panic(status, "out of memory");
    if (r.data) std::memset(r.data, 0, size_t(size));     // RustBuffer::new_with_size: vec![0u8; size], len == capacity
    return r;
}
RustBuffer ffi_slimg_ffi_rustbuffer_from_bytes(ForeignBytes bytes, RustCallStatus *status) {
    set_ok(status);
    if (bytes.len < 0 || (bytes.len > 0 && !bytes.data)) return panic(status, "bad ForeignBytes");
    return make_buffer(bytes.data, size_t(bytes.len));
}
void ffi_slimg_ffi_rustbuffer_free(RustBuffer buf, RustCallStatus *status) {
    set_ok(status);
    std::free(buf.data);
}
RustBuffer ffi_slimg_ffi_rustbuffer_reserve(RustBuffer buf, uint64_t additional, RustCallStatus *status) {
    set_ok(status);
    const uint64_t want = buf.len + additional;
    if (want <= buf.capacity) return buf;
    uint8_t *p = static_cast<uint8_t *>(std::realloc(buf.data, size_t(want)));
    if (!p) return panic(status, "out of memory");
    buf.data = p; buf.capacity = want;
    return buf;
}

// ---- contract ------------------------------------------------------------------------------------------------
uint32_t ffi_slimg_ffi_uniffi_contract_version(void) { return 30; }                 // slimg_ffi.kt:809
// values the checked-in Kotlin binding verifies at load (slimg_ffi.kt:818-852)
uint16_t uniffi_slimg_ffi_checksum_func_convert(void) { return 14232; }
uint16_t uniffi_slimg_ffi_checksum_func_crop(void) { return 14284; }
uint16_t uniffi_slimg_ffi_checksum_func_decode(void) { return 62226; }
uint16_t uniffi_slimg_ffi_checksum_func_decode_file(void) { return 64209; }
uint16_t uniffi_slimg_ffi_checksum_func_extend(void) { return 52877; }
uint16_t uniffi_slimg_ffi_checksum_func_format_can_encode(void) { return 31855; }
uint16_t uniffi_slimg_ffi_checksum_func_format_extension(void) { return 23550; }
uint16_t uniffi_slimg_ffi_checksum_func_format_from_extension(void) { return 56933; }
uint16_t uniffi_slimg_ffi_checksum_func_format_from_magic_bytes(void) { return 41284; }
uint16_t uniffi_slimg_ffi_checksum_func_optimize(void) { return 28797; }
uint16_t uniffi_slimg_ffi_checksum_func_output_path(void) { return 60410; }
uint16_t uniffi_slimg_ffi_checksum_func_resize(void) { return 48866; }

void slimg_b200_set_codec_hooks(slimg_encode_hook enc, slimg_decode_hook dec, void *user) {
    g_hooks.enc = enc; g_hooks.dec = dec; g_hooks.user = user;
}
void slimg_b200_set_jpeg_coefficient_hook(slimg_jpeg_coef_hook hook, void *user, int32_t table_base, uint32_t flags) {
    g_hooks.jpeg = hook; g_hooks.jpeg_user = user; g_hooks.jpeg_tables = table_base; g_hooks.jpeg_flags = flags;
}

// ---- pixel functions -----------------------------------------------------------------------------------------
RustBuffer uniffi_slimg_ffi_fn_func_resize(RustBuffer image, RustBuffer mode, RustCallStatus *status) {   // lib.rs:334-339
    Reader ri(image), rm(mode);
    slimg_image in;
    slimg_resize_mode m;
    RustBuffer ret = kEmpty;
    if (!read_image(ri, in) || !ri.done() || !read_resize(rm, m) || !rm.done()) ret = panic(status, "resize: malformed argument buffer");
    else {
        slimg_image out{0, 0, nullptr, 0};
        const slimg_status s = slimg_b200_resize(&in, &m, &out);
        if (s != SLIMG_OK) ret = raise_last(status, s);
        else { ret = lower_image(out); slimg_b200_image_free(&out); set_ok(status); }
    }
    release(image); release(mode);
    return ret;
}

RustBuffer uniffi_slimg_ffi_fn_func_crop(RustBuffer image, RustBuffer mode, RustCallStatus *status) {     // lib.rs:320-325
    Reader ri(image), rm(mode);
    slimg_image in;
    slimg_crop_mode m;
    RustBuffer ret = kEmpty;
    if (!read_image(ri, in) || !ri.done() || !read_crop(rm, m) || !rm.done()) ret = panic(status, "crop: malformed argument buffer");
    else {
        slimg_image out{0, 0, nullptr, 0};
        const slimg_status s = slimg_b200_crop(&in, &m, &out);
        if (s != SLIMG_OK) ret = raise_last(status, s);
        else { ret = lower_image(out); slimg_b200_image_free(&out); set_ok(status); }
    }
    release(image); release(mode);
    return ret;
}

RustBuffer uniffi_slimg_ffi_fn_func_extend(RustBuffer image, RustBuffer mode, RustBuffer fill, RustCallStatus *status) {   // lib.rs:327-332
    Reader ri(image), rm(mode), rf(fill);
    slimg_image in;
    slimg_extend_mode m;
    slimg_fill_color f;
    RustBuffer ret = kEmpty;
    if (!read_image(ri, in) || !ri.done() || !read_extend(rm, m) || !rm.done() || !read_fill(rf, f) || !rf.done())
        ret = panic(status, "extend: malformed argument buffer");
    else {
        slimg_image out{0, 0, nullptr, 0};
        const slimg_status s = slimg_b200_extend(&in, &m, &f, &out);
        if (s != SLIMG_OK) ret = raise_last(status, s);
        else { ret = lower_image(out); slimg_b200_image_free(&out); set_ok(status); }
    }
    release(image); release(mode); release(fill);
    return ret;
}

RustBuffer uniffi_slimg_ffi_fn_func_convert(RustBuffer image, RustBuffer options, RustCallStatus *status) {   // lib.rs:297-318
    Reader ri(image), ro(options);
    slimg_image in;
    slimg_pipeline_options o{};
    int32_t format = 0;
    bool ok = read_image(ri, in) && ri.done() && read_format(ro, format);
    const uint8_t quality = ro.u8();
    bool has = false;
    if (ok && (ok = read_option_tag(ro, has)) && has) { o.has_resize = 1; ok = read_resize(ro, o.resize); }
    if (ok && (ok = read_option_tag(ro, has)) && has) { o.has_crop = 1; ok = read_crop(ro, o.crop); }
    if (ok && (ok = read_option_tag(ro, has)) && has) { o.has_extend = 1; ok = read_extend(ro, o.extend); }
    if (ok && (ok = read_option_tag(ro, has)) && has) { o.has_fill = 1; ok = read_fill(ro, o.fill); }
    RustBuffer ret = kEmpty;
    if (!ok || !ro.done()) ret = panic(status, "convert: malformed argument buffer");
    else {
        if (format == 1 && g_hooks.jpeg) {
            // pipeline.rs:71-95 with codec/jpeg.rs:51-72 folded in: pixel stages and the JPEG front end run back to
            // back on the device, only coefficients come back, the installed entropy coder writes the file
            uint16_t ql[64], qc[64];
            slimg_b200_jpeg_quant_table(quality, 0, g_hooks.jpeg_tables, ql);       // codec/jpeg.rs:60 set_quality
            slimg_b200_jpeg_quant_table(quality, 1, g_hooks.jpeg_tables, qc);
            slimg_jpeg_coefs coefs;
            const slimg_status s = slimg_b200_pipeline_jpeg(&in, &o, SLIMG_ORDER_CONVERT, ql, qc, g_hooks.jpeg_flags, &coefs);
            if (s != SLIMG_OK) ret = raise_last(status, s);
            else {
                uint8_t *out = nullptr;
                size_t out_len = 0;
                const char *msg = nullptr;
                const int32_t rc = g_hooks.jpeg(g_hooks.jpeg_user, &coefs, quality, &out, &out_len, &msg);
                if (rc != SLIMG_OK) ret = raise(status, rc, msg ? msg : "JPEG entropy coder hook failed");
                else { ret = lower_pipeline_result(out, out_len, format, coefs.geom.width, coefs.geom.height); set_ok(status); }
                std::free(out);
                slimg_b200_jpeg_coefs_free(&coefs);
            }
        } else {
            slimg_image px{0, 0, nullptr, 0};
            const slimg_status s = slimg_b200_pipeline(&in, &o, SLIMG_ORDER_CONVERT, &px);      // pipeline.rs:71-89
            if (s != SLIMG_OK) ret = raise_last(status, s);
            else { ret = encode_with_hook(status, px, format, quality); slimg_b200_image_free(&px); }
        }
    }
    release(image); release(options);
    return ret;
}

// ---- codec functions: hooks ----------------------------------------------------------------------------------
static RustBuffer decode_bytes(const uint8_t *d, size_t n, RustCallStatus *status) {        // lib.rs:277-284
    if (!g_hooks.dec) return raise(status, SLIMG_ERR_DECODE, "no decoder linked: decoding stays on the reference codecs (slimg_b200_set_codec_hooks)");
    slimg_image px{0, 0, nullptr, 0};
    int32_t format = 0;
    const char *msg = nullptr;
    const int32_t rc = g_hooks.dec(g_hooks.user, d, n, &px, &format, &msg);
    if (rc != SLIMG_OK) { std::free(px.data); return raise(status, rc, msg ? msg : "decoder hook failed"); }
    Writer w;
    w.b.reserve(16 + px.len);
    w.u32(px.width); w.u32(px.height); w.i32(int32_t(px.len)); w.bytes(px.data, px.len); w.i32(format);   // DecodeResult
    std::free(px.data);
    set_ok(status);
    return take(w);
}

RustBuffer uniffi_slimg_ffi_fn_func_decode(RustBuffer data, RustCallStatus *status) {
    Reader r(data);
    const int32_t n = r.i32();
    const uint8_t *d = n >= 0 ? r.bytes(size_t(n)) : nullptr;
    RustBuffer ret = (!r.ok || n < 0 || !r.done()) ? panic(status, "decode: malformed argument buffer") : decode_bytes(d, size_t(n), status);
    release(data);
    return ret;
}

RustBuffer uniffi_slimg_ffi_fn_func_decode_file(RustBuffer path, RustCallStatus *status) {   // lib.rs:287-294; a top-level String is raw UTF-8
    const std::string p(reinterpret_cast<const char *>(path.data), size_t(path.len));
    release(path);
    FILE *f = std::fopen(p.c_str(), "rb");
    if (!f) return raise(status, SLIMG_ERR_IO, "cannot open " + p);
    std::vector<uint8_t> buf;
    uint8_t chunk[65536];
    size_t k;
    while ((k = std::fread(chunk, 1, sizeof chunk, f)) > 0) buf.insert(buf.end(), chunk, chunk + k);
    std::fclose(f);
    return decode_bytes(buf.data(), buf.size(), status);
}

RustBuffer uniffi_slimg_ffi_fn_func_optimize(RustBuffer data, uint8_t quality, RustCallStatus *status) {   // lib.rs:342-351
    Reader r(data);
    const int32_t n = r.i32();
    const uint8_t *d = n >= 0 ? r.bytes(size_t(n)) : nullptr;
    RustBuffer ret = kEmpty;
    if (!r.ok || n < 0 || !r.done()) ret = panic(status, "optimize: malformed argument buffer");
    else if (!g_hooks.dec) ret = raise(status, SLIMG_ERR_DECODE, "no decoder linked: decoding stays on the reference codecs (slimg_b200_set_codec_hooks)");
    else {
        slimg_image px{0, 0, nullptr, 0};
        int32_t format = 0;
        const char *msg = nullptr;
        const int32_t rc = g_hooks.dec(g_hooks.user, d, size_t(n), &px, &format, &msg);
        if (rc != SLIMG_OK) ret = raise(status, rc, msg ? msg : "decoder hook failed");
        else ret = encode_with_hook(status, px, format, quality);
        std::free(px.data);
    }
    release(data);
    return ret;
}

// ---- host-only helpers ---------------------------------------------------------------------------------------
RustBuffer uniffi_slimg_ffi_fn_func_format_extension(RustBuffer format, RustCallStatus *status) {   // lib.rs:254-257
    Reader r(format);
    int32_t f = 0;
    const bool ok = read_format(r, f) && r.done();
    release(format);
    if (!ok) return panic(status, "format_extension: malformed argument buffer");
    set_ok(status);
    return make_buffer(kExt[f], std::strlen(kExt[f]));            // a top-level String is raw UTF-8
}

int8_t uniffi_slimg_ffi_fn_func_format_can_encode(RustBuffer format, RustCallStatus *status) {     // lib.rs:260-263; format.rs:73-75
    Reader r(format);
    int32_t f = 0;
    const bool ok = read_format(r, f) && r.done();
    release(format);
    if (!ok) { panic(status, "format_can_encode: malformed argument buffer"); return 0; }
    set_ok(status);
    return 1;
}

static RustBuffer lower_option_format(int32_t f) {
    Writer w;
    if (f) { w.u8(1); w.i32(f); } else w.u8(0);
    return take(w);
}

RustBuffer uniffi_slimg_ffi_fn_func_format_from_extension(RustBuffer path, RustCallStatus *status) {   // lib.rs:266-269
    const std::string p(reinterpret_cast<const char *>(path.data), size_t(path.len));
    release(path);
    set_ok(status);
    return lower_option_format(format_from_ext(p));
}

RustBuffer uniffi_slimg_ffi_fn_func_format_from_magic_bytes(RustBuffer data, RustCallStatus *status) {   // lib.rs:272-275
    Reader r(data);
    const int32_t n = r.i32();
    const uint8_t *d = n >= 0 ? r.bytes(size_t(n)) : nullptr;
    RustBuffer ret = (!r.ok || n < 0 || !r.done()) ? panic(status, "format_from_magic_bytes: malformed argument buffer")
                                                     : (set_ok(status), lower_option_format(format_from_magic(d, size_t(n))));
    release(data);
    return ret;
}

RustBuffer uniffi_slimg_ffi_fn_func_output_path(RustBuffer input, RustBuffer format, RustBuffer output, RustCallStatus *status) {   // lib.rs:354-362
    const std::string in(reinterpret_cast<const char *>(input.data), size_t(input.len));
    Reader rf(format), ro(output);
    int32_t f = 0;
    bool ok = read_format(rf, f) && rf.done();
    bool has = false;
    std::string out;
    if (ok && (ok = read_option_tag(ro, has)) && has) {
        const int32_t n = ro.i32();
        const uint8_t *d = n >= 0 ? ro.bytes(size_t(n)) : nullptr;
        ok = ro.ok && n >= 0;
        if (ok) out.assign(reinterpret_cast<const char *>(d), size_t(n));
    }
    ok = ok && ro.done();
    release(input); release(format); release(output);
    if (!ok) return panic(status, "output_path: malformed argument buffer");
    std::string res;                                               // pipeline.rs:130-141
    struct stat sb;
    if (!has) res = with_extension(in, kExt[f]);
    else if (::stat(out.c_str(), &sb) == 0 && S_ISDIR(sb.st_mode)) {
        std::string dir = out;
        if (!dir.empty() && dir.back() != '/') dir += '/';
        res = with_extension(dir + file_stem(in), kExt[f]);
    } else res = out;
    set_ok(status);
    return make_buffer(res.data(), res.size());
}

}  // extern "C"
