// Host-side Wavefront .obj reader with the reference's grammar (scene_reader.py:29-46 line cleaning,
// :49-104 `v` / `f` / fan triangulation / negative indices / "Skipping command"), written to load the
// 10^6..10^7-triangle meshes of configs 4 and 5 in seconds: the file is read once, split at line ends
// into one slice per host thread, and numbers go through std::from_chars (correctly rounded, so the
// doubles are bit-identical to Python's float()).  No CUDA in here.
#include <thread>
#include <string>
#include <cerrno>
#include <charconv>

namespace ptbobj {

static int fail(int code, const std::string &msg) { set_error("%s", msg.c_str()); return code; }

struct Slice {
    const char *lo, *hi;
    std::vector<double> v;            // xyz per vertex
    std::vector<int64_t> f;           // triangles; relative (negative) indices are stored as i - REL_BIAS
    std::vector<int64_t> f_nv;        // per triangle corner: slice-local vertex count when the face was read
    std::string skipped;
    int err = 0;
};

static const int64_t REL_BIAS = (int64_t)1 << 40;     // an absolute index is >= -1 ('f 0 ..' gives -1, as in the reference)

// float(token): false = not a number the reference would accept.  std::from_chars is correctly rounded
// (as Python's float() is), locale-free and takes ~25 ns for a 17-digit repr.
static bool parse_double(const char *s, const char *e, double *out) {
    const char *p = s;
    if (p < e && *p == '+') { p++; if (p < e && (*p == '-' || *p == '+')) return false; }
    if (p == e) return false;
    double d;
    auto r = std::from_chars(p, e, d, std::chars_format::general);
    if (r.ec == std::errc() && r.ptr == e) { *out = d; return true; }
    if (r.ec != std::errc::result_out_of_range || r.ptr != e) return false;
    // overflow / underflow: Python gives +-inf / 0 / a denormal, as strtod does
    size_t n = (size_t)(e - s);
    if (n > 400) return false;
    char buf[408];
    memcpy(buf, s, n); buf[n] = 0;
    char *end = nullptr;
    d = strtod(buf, &end);
    if (end != buf + n) return false;
    *out = d;
    return true;
}

static bool parse_int(const char *s, const char *e, int64_t *out) {          // int(token)
    const char *p = s;
    if (p < e && *p == '+') { p++; if (p < e && (*p == '-' || *p == '+')) return false; }
    if (p == e) return false;
    long long v;
    auto r = std::from_chars(p, e, v, 10);
    if (r.ec != std::errc() || r.ptr != e) return false;
    *out = v;
    return true;
}

static void parse_slice(Slice *sl) {
    const char *p = sl->lo;
    std::vector<std::pair<const char *, const char *>> tok;
    std::vector<int64_t> face;
    while (p < sl->hi) {
        const char *eol = p;
        while (eol < sl->hi && *eol != '\n' && *eol != '\r') eol++;
        const char *end = p;
        while (end < eol && *end != '#') end++;            // leading-'#' lines and trailing comments alike
        tok.clear();
        const char *q = p;
        while (q < end) {
            while (q < end && (*q == ' ' || *q == '\t')) q++;
            const char *t0 = q;
            while (q < end && *q != ' ' && *q != '\t') q++;
            if (q > t0) tok.emplace_back(t0, q);
        }
        p = eol < sl->hi ? eol + 1 : eol;
        if (tok.empty()) continue;
        size_t l0 = (size_t)(tok[0].second - tok[0].first);
        if (l0 == 1 && *tok[0].first == 'v') {
            double xyz[3];
            // the reference keeps however many coordinates the line has and only objects when a face mixes them
            // (vector.py:63-72): anything but x y z is left to the caller's line-by-line parser
            if (tok.size() != 4) { sl->err = PTB_ERR_PARSE_IRREGULAR; return; }
            for (size_t i = 1; i < tok.size(); i++) {
                double d;
                if (!parse_double(tok[i].first, tok[i].second, &d)) { sl->err = PTB_ERR_PARSE; return; }
                if (i <= 3) xyz[i - 1] = d;
            }
            sl->v.insert(sl->v.end(), xyz, xyz + 3);
        } else if (l0 == 1 && *tok[0].first == 'f') {
            face.clear();
            for (size_t i = 1; i < tok.size(); i++) {
                int64_t k;
                if (!parse_int(tok[i].first, tok[i].second, &k)) { sl->err = PTB_ERR_PARSE; return; }   // 'i/j/k' too
                if (k <= -REL_BIAS / 2 || k >= REL_BIAS) { sl->err = PTB_ERR_PARSE; return; }
                face.push_back(k < 0 ? k - REL_BIAS : k - 1);
            }
            if (face.size() < 3) { sl->err = PTB_ERR_PARSE_IRREGULAR; return; }
            const int64_t nv = (int64_t)(sl->v.size() / 3);
            for (size_t i = 1; i + 1 < face.size(); i++) {
                const int64_t tri[3] = {face[0], face[i], face[i + 1]};
                sl->f.insert(sl->f.end(), tri, tri + 3);
                sl->f_nv.push_back(nv);
            }
        } else {
            sl->skipped.append(tok[0].first, (size_t)(end - tok[0].first));
            sl->skipped.push_back('\n');
        }
    }
}

}  // namespace ptbobj

extern "C" int ptb_obj_parse(const char *path, double **verts, int64_t *n_verts, int64_t **faces, int64_t *n_faces,
                             char **skipped, int threads) {
    using namespace ptbobj;
    if (!path || !verts || !n_verts || !faces || !n_faces) return fail(PTB_ERR_ARG, "ptb_obj_parse: null argument");
    *verts = nullptr; *faces = nullptr; *n_verts = *n_faces = 0;
    if (skipped) *skipped = nullptr;
    FILE *fp = fopen(path, "rb");
    if (!fp) return fail(PTB_ERR_IO, std::string("ptb_obj_parse: cannot open ") + path);
    fseek(fp, 0, SEEK_END);
    const long size = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    std::vector<char> text((size_t)(size > 0 ? size : 0));
    if (size > 0 && fread(text.data(), 1, (size_t)size, fp) != (size_t)size) { fclose(fp); return fail(PTB_ERR_IO, "ptb_obj_parse: short read"); }
    fclose(fp);
    const char *lo = text.data(), *hi = lo + text.size();
    // the reference's remove_spaces_from_start('') : a final line of blanks without a newline (scene_reader.py:31)
    if (hi > lo && hi[-1] == ' ') {
        const char *q = hi;
        while (q > lo && q[-1] == ' ') q--;
        if (q == lo || q[-1] == '\n' || q[-1] == '\r') return fail(PTB_ERR_PARSE_INDEX, "ptb_obj_parse: blank last line");
    }
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min(nt, 64));
    if (text.size() < ((size_t)1 << 20)) nt = 1;
    std::vector<Slice> sl((size_t)nt);
    const char *cur = lo;
    for (int i = 0; i < nt; i++) {
        const char *cut = i == nt - 1 ? hi : lo + text.size() * (size_t)(i + 1) / (size_t)nt;
        if (cut < cur) cut = cur;
        while (cut < hi && (cut == lo || cut[-1] != '\n')) cut++;      // advance to just after a newline
        sl[i].lo = cur; sl[i].hi = cut;
        cur = cut;
    }
    if (nt == 1) parse_slice(&sl[0]);
    else {
        std::vector<std::thread> th;
        for (int i = 0; i < nt; i++) th.emplace_back(parse_slice, &sl[i]);
        for (auto &t : th) t.join();
    }
    int64_t nv = 0, nf = 0;
    size_t nskip = 0;
    for (auto &s : sl) {
        if (s.err) return fail(s.err, (s.err == PTB_ERR_PARSE_IRREGULAR ? "ptb_obj_parse: a 'v' line without exactly 3 coordinates or an 'f' line with fewer than 3 indices in "
                                                                         : "ptb_obj_parse: malformed 'v' or 'f' line in ") + std::string(path));
        nv += (int64_t)(s.v.size() / 3); nf += (int64_t)(s.f.size() / 3); nskip += s.skipped.size();
    }
    double *V = (double *)malloc(sizeof(double) * (size_t)std::max<int64_t>(nv * 3, 1));
    int64_t *F = (int64_t *)malloc(sizeof(int64_t) * (size_t)std::max<int64_t>(nf * 3, 1));
    char *S = skipped ? (char *)malloc(nskip + 1) : nullptr;
    if (!V || !F || (skipped && !S)) { free(V); free(F); free(S); return fail(PTB_ERR_NOMEM, "ptb_obj_parse: out of host memory"); }
    std::atomic<int> irregular{0};
    auto stitch = [&](int i, int64_t v0, int64_t f0) {
        Slice &s = sl[(size_t)i];
        if (!s.v.empty()) memcpy(V + v0 * 3, s.v.data(), sizeof(double) * s.v.size());
        for (size_t t = 0; t < s.f_nv.size(); t++)
            for (int c = 0; c < 3; c++) {
                int64_t k = s.f[t * 3 + c];
                const int64_t nv_now = v0 + s.f_nv[t];                        // vertices read so far (scene_reader.py:71-72)
                if (k < -REL_BIAS / 2) k = nv_now + (k + REL_BIAS);           // nv + i
                // the reference looks the index up among the vertices read SO FAR, with Python's wrap-around for
                // negatives and IndexError beyond: only plain in-range indices are resolved here
                if (k < 0 || k >= nv_now) irregular.store(1, std::memory_order_relaxed);
                F[(f0 + (int64_t)t) * 3 + c] = k;
            }
    };
    {
        std::vector<std::thread> th;
        int64_t v0 = 0, f0 = 0;
        size_t s0 = 0;
        for (int i = 0; i < nt; i++) {
            if (nt == 1) stitch(i, v0, f0); else th.emplace_back(stitch, i, v0, f0);
            if (S) { memcpy(S + s0, sl[(size_t)i].skipped.data(), sl[(size_t)i].skipped.size()); s0 += sl[(size_t)i].skipped.size(); }
            v0 += (int64_t)(sl[(size_t)i].v.size() / 3); f0 += (int64_t)(sl[(size_t)i].f.size() / 3);
        }
        for (auto &t : th) t.join();
        if (S) S[s0] = 0;
    }
    if (irregular.load()) {
        free(V); free(F); free(S);
        return fail(PTB_ERR_PARSE_IRREGULAR, "ptb_obj_parse: a face index lies outside the vertices read before it in " + std::string(path));
    }
    *verts = V; *n_verts = nv; *faces = F; *n_faces = nf;
    if (skipped) *skipped = S;
    return PTB_OK;
}

extern "C" void ptb_obj_free(void *p) { free(p); }

// Writer for generated meshes (synth.write_obj): "v x y z" lines with the shortest decimals that read back to the same
// doubles (std::to_chars, what Python's repr() prints up to notation), then "f i j k" lines, 1-based.
extern "C" int ptb_obj_write(const char *path, const double *verts, int64_t n_verts, const int64_t *faces, int64_t n_faces,
                             const char *header) {
    if (!path || (n_verts && !verts) || (n_faces && !faces)) { set_error("ptb_obj_write: null argument"); return PTB_ERR_ARG; }
    FILE *fp = fopen(path, "wb");
    if (!fp) { set_error("ptb_obj_write: cannot open %s", path); return PTB_ERR_IO; }
    std::vector<char> buf;
    buf.reserve(1 << 22);
    auto flush = [&]() { bool ok = buf.empty() || fwrite(buf.data(), 1, buf.size(), fp) == buf.size(); buf.clear(); return ok; };
    bool ok = true;
    if (header && *header) { buf.push_back('#'); buf.push_back(' '); buf.insert(buf.end(), header, header + strlen(header)); buf.push_back('\n'); }
    char tmp[40];
    for (int64_t i = 0; i < n_verts && ok; ++i) {
        buf.push_back('v');
        for (int k = 0; k < 3; ++k) {
            buf.push_back(' ');
            const double x = verts[3 * i + k];
            auto r = std::to_chars(tmp, tmp + sizeof(tmp), x);
            buf.insert(buf.end(), tmp, r.ptr);
        }
        buf.push_back('\n');
        if (buf.size() > (1u << 22) - 256) ok = flush();
    }
    for (int64_t i = 0; i < n_faces && ok; ++i) {
        buf.push_back('f');
        for (int k = 0; k < 3; ++k) {
            buf.push_back(' ');
            auto r = std::to_chars(tmp, tmp + sizeof(tmp), (long long)(faces[3 * i + k] + 1));
            buf.insert(buf.end(), tmp, r.ptr);
        }
        buf.push_back('\n');
        if (buf.size() > (1u << 22) - 256) ok = flush();
    }
    ok = flush() && ok;
    ok = fclose(fp) == 0 && ok;
    if (!ok) { set_error("ptb_obj_write: write to %s failed", path); return PTB_ERR_IO; }
    return PTB_OK;
}
