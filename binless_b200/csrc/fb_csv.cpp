// fb_csv.cpp -- the reference's csvfile as built (csvfile.hpp:30-117 with USE_BOOST = 1, binless/src/Makevars:1-2):
// a gzip stream (boost::iostreams::gzip_compressor, csvfile.hpp:46-48) of a quoted header row and one row per cell,
// every value printed as std::fixed << std::setprecision(4), comma separated, rows ended by std::endl
// (fast_dataframe.cpp:23-57). chromosome.R:232-241 names the file *_tmp.csv.gz and reads it back with read.csv.
//
// Rows are formatted AND deflated by several host threads, slab by slab: every thread turns its rows into text and
// compresses the text as a raw deflate block sequence that ends on a byte boundary (Z_SYNC_FLUSH; the last one of the
// file with Z_FINISH), so the pieces concatenate into ONE deflate stream = one gzip member, as the reference writes.
// The CRC-32 of the whole text is combined from the per-piece CRCs (crc32_combine).
#include "fb_csv.hpp"
#include <zlib.h>
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <string>
#include <thread>
#include <vector>

namespace fb {

namespace {

// Compression level of the gzip file. The reference writes zlib's default (6); on these rows (digits that hardly repeat)
// level 6 deflates 7.6 MB/s per thread for 0.45 of the text, level 1 46 MB/s for 0.50, and the file is a temporary that
// chromosome.R reads straight back: level 1 unless FB_CSV_LEVEL (0-9) says otherwise. The text inside is the same.
int deflate_level() {
    static const int v = [] {
        const char *e = std::getenv("FB_CSV_LEVEL");
        return e ? std::max(0, std::min(9, std::atoi(e))) : 1;
    }();
    return v;
}

struct Piece { std::string z; uLong crc = 0; size_t len = 0; bool ok = true; };

// raw deflate of `text`, ending on a byte boundary (or, when `last`, with the final block)
void deflate_piece(const std::string &text, bool last, Piece &out) {
    out.len = text.size();
    out.crc = crc32(crc32(0L, Z_NULL, 0), reinterpret_cast<const Bytef *>(text.data()), (uInt)text.size());
    z_stream zs;
    std::memset(&zs, 0, sizeof zs);
    if (deflateInit2(&zs, deflate_level(), Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) { out.ok = false; return; }
    out.z.resize(deflateBound(&zs, (uLong)text.size()) + 16);
    zs.next_in = reinterpret_cast<Bytef *>(const_cast<char *>(text.data()));
    zs.avail_in = (uInt)text.size();
    zs.next_out = reinterpret_cast<Bytef *>(&out.z[0]);
    zs.avail_out = (uInt)out.z.size();
    const int rc = deflate(&zs, last ? Z_FINISH : Z_SYNC_FLUSH);
    out.ok = last ? rc == Z_STREAM_END : (rc == Z_OK && zs.avail_in == 0);
    out.z.resize(out.z.size() - zs.avail_out);
    deflateEnd(&zs);
}

// "%.4f" of a double without printf (glibc's %f goes through multi-precision arithmetic: ~850 ns per value and thread,
// the whole cost of a plain-text file and half of a gzip one). Exact: |x| = m 2^(e-52) with a 53-bit integer m, so
// |x| 10^4 = (10^4 m) 2^-(52-e) is a 67-bit integer shifted right; the dropped bits decide the rounding, ties to even,
// which is what printf and std::fixed << std::setprecision(4) produce in the default rounding mode. Returns the number
// of characters written (at most 24), or 0 for what is left to snprintf: NaN, infinities, |x| >= 2^50.
int fmt_fixed4(char *out, double x) {
    uint64_t bits;
    std::memcpy(&bits, &x, sizeof bits);
    const int be = (int)((bits >> 52) & 0x7ff);
    if (be >= 1023 + 50) return 0;
    uint64_t m = bits & ((uint64_t(1) << 52) - 1);
    int e = be - 1023;
    if (be == 0) e = -1022; else m |= uint64_t(1) << 52;       // subnormals (and zero) have no implicit bit
    const unsigned __int128 p = (unsigned __int128)m * 10000u;
    const int s = 52 - e;                                       // >= 3
    uint64_t q;
    if (s >= 100) q = 0;                                        // p < 2^67: far below one half
    else {
        q = (uint64_t)(p >> s);
        const unsigned __int128 rem = p & (((unsigned __int128)1 << s) - 1), half = (unsigned __int128)1 << (s - 1);
        if (rem > half || (rem == half && (q & 1u))) ++q;
    }
    char *o = out;
    if (bits >> 63) *o++ = '-';
    uint64_t ip = q / 10000u;
    const unsigned fp = (unsigned)(q % 10000u);
    char tmp[20];
    int n = 0;
    do { tmp[n++] = (char)('0' + ip % 10u); ip /= 10u; } while (ip);
    while (n) *o++ = tmp[--n];
    *o++ = '.';
    *o++ = (char)('0' + fp / 1000u); *o++ = (char)('0' + fp / 100u % 10u); *o++ = (char)('0' + fp / 10u % 10u); *o++ = (char)('0' + fp % 10u);
    return (int)(o - out);
}

void put_le32(unsigned char *p, uLong v) { p[0] = v & 255; p[1] = (v >> 8) & 255; p[2] = (v >> 16) & 255; p[3] = (v >> 24) & 255; }

} // namespace

bool write_csv_gz(const char *path, const char *header, int64_t nrows, int ncols, const double *const *cols, bool gzip) {
    FILE *f = std::fopen(path, "wb");
    if (!f) { std::printf("Exception creating csv file: cannot open %s\n", path); return false; } // the reference only reports it
    bool ok = true;
    int nthr = (int)std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
    if (const char *e = std::getenv("FB_HOST_THREADS")) nthr = std::max(1, std::min(64, std::atoi(e)));
    const int64_t slab = int64_t(1) << 20;      // rows per slab: a few tens of MB of text in flight
    uLong crc = crc32(0L, Z_NULL, 0);
    uint64_t total = 0;
    if (gzip) {
        // gzip member header (RFC 1952): no name, no time stamp, as boost's gzip_compressor writes by default
        const unsigned char hdr[10] = {0x1f, 0x8b, 8, 0, 0, 0, 0, 0, 0, 3 /* OS = Unix */};
        ok &= std::fwrite(hdr, 1, sizeof hdr, f) == sizeof hdr;
    }
    std::vector<std::string> text(nthr);
    std::vector<Piece> piece(nthr);
    const int64_t nslab = nrows > 0 ? (nrows + slab - 1) / slab : 1;
    for (int64_t s = 0; s < nslab && ok; ++s) {
        const int64_t base = s * slab, cnt = std::max<int64_t>(0, std::min<int64_t>(slab, nrows - base));
        const bool last_slab = s + 1 == nslab;
        std::vector<std::future<void>> fut;
        for (int t = 0; t < nthr; ++t)
            fut.push_back(std::async(std::launch::async, [&, t] {
                const int64_t lo = base + cnt * t / nthr, hi = base + cnt * (t + 1) / nthr;
                std::string &o = text[t];
                o.clear(); o.reserve((size_t)(hi - lo) * 10 * ncols + 256);
                if (s == 0 && t == 0) { o.append(header); o.push_back('\n'); }
                char buf[64];
                for (int64_t i = lo; i < hi; ++i)
                    for (int c = 0; c < ncols; ++c) {
                        int len = fmt_fixed4(buf, cols[c][i]);
                        if (!len) len = std::snprintf(buf, sizeof buf, "%.4f", cols[c][i]);
                        if (len < (int)sizeof buf) o.append(buf, (size_t)len);
                        else { std::string big((size_t)len + 1, '\0'); std::snprintf(&big[0], big.size(), "%.4f", cols[c][i]); o.append(big.c_str(), (size_t)len); }
                        o.push_back(c + 1 == ncols ? '\n' : ',');
                    }
                if (gzip) deflate_piece(o, last_slab && t == nthr - 1, piece[t]);
            }));
        for (auto &x : fut) x.get();
        for (int t = 0; t < nthr && ok; ++t) {
            if (gzip) {
                ok &= piece[t].ok && std::fwrite(piece[t].z.data(), 1, piece[t].z.size(), f) == piece[t].z.size();
                crc = crc32_combine(crc, piece[t].crc, (z_off_t)piece[t].len);
                total += piece[t].len;
            } else ok &= std::fwrite(text[t].data(), 1, text[t].size(), f) == text[t].size();
        }
    }
    if (gzip && ok) {
        unsigned char tr[8];
        put_le32(tr, crc); put_le32(tr + 4, (uLong)(total & 0xffffffffu));
        ok &= std::fwrite(tr, 1, 8, f) == 8;
    }
    ok &= std::fclose(f) == 0;
    if (!ok) std::printf("Exception creating csv file: write to %s failed\n", path);
    return ok;
}

} // namespace fb
