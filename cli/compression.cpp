// Linux `Compression` front end on the C ABI: the -c / -d / -n modes of the reference's
// Compression/Compression.cpp:96-182 and the matching-pursuit half of its -g sweep (:303-357), same arguments, same lines
// on stdout, the GPU library behind them.
//
//   Compression -c <quality 1.0-8.0 | max> <input.ppm> <output.mn>        Compression.cpp:102-133
//   Compression -d <input.mn> <output.ppm>                                 Compression.cpp:134-146
//   Compression -n <quality 1.0-8.0 | max> <input.ppm> [output.ppm]        Compression.cpp:147-182
//   Compression -g <input folder> <output statistics.csv>                  Compression.cpp:303-357 (MP rows only)
//
// Differences, all forced by the platform: images are binary PPM (P6) — the reference loads and saves through GDI+
// (ImageHelper/src/imgloader.cpp), which does not exist here; its own PNM code path (ImageHelper/src/pnmfile.cpp) is the
// format kept. The dictionary is imported data (MNENC_DICT, default imageexperiments_b200/data/dict_bs8.mndict next to the
// library), not rebuilt per run. -g walks the folder for .ppm files (the reference: .jpg / .tif) in sorted
// order and writes the "MP" rows of the reference's CSV; its "JPEG" rows, -j and -s (JPEG comparison through GDI+, training
// statistics of the double-precision twin) are out of scope.
// There is no CPU path: without a B200 every mode fails with the library's error text.
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <filesystem>
#include <format>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "../include/mnenc.h"

namespace {

constexpr int kBlockSize = 8;  // Compression.cpp:28
constexpr int kK = 32;         // Compression.cpp:29

struct Image {
    size_t w = 0, h = 0;
    std::vector<uint8_t> px;  // packed RGB8, row-major (img::image<rgb>::data)
};

[[noreturn]] void die(const std::string& what) {
    std::cerr << "Compression: " << what << std::endl;
    std::exit(1);
}
[[noreturn]] void dieLib(const char* call) { die(std::string(call) + ": " + mnenc_last_error()); }

// Binary PPM, comments allowed between header fields (pnmfile.cpp:45-60,142-170).
Image loadPpm(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) die("cannot open " + path);
    auto field = [&]() {
        std::string tok;
        for (;;) {
            int c = f.peek();
            if (c == '#') {
                std::string line;
                std::getline(f, line);
            } else if (c == ' ' || c == '\t' || c == '\n' || c == '\r') {
                f.get();
            } else {
                break;
            }
        }
        f >> tok;
        return tok;
    };
    if (field() != "P6") die(path + ": not a binary PPM (P6)");
    Image im;
    auto number = [&](long limit) {  // a decimal field with nothing after the digits, 1 .. limit
        const std::string tok = field();
        if (tok.empty() || tok.size() > 7 || tok.find_first_not_of("0123456789") != std::string::npos) die(path + ": malformed PPM header");
        const long v = std::atol(tok.c_str());
        if (v <= 0 || v > limit) die(path + ": unsupported PPM header (need 8-bit RGB, sides up to 2^20)");
        return v;
    };
    const long w = number(1L << 20), h = number(1L << 20), maxv = number(255);
    if (maxv != 255) die(path + ": unsupported PPM header (need 8-bit RGB)");
    if (static_cast<unsigned long long>(w) * static_cast<unsigned long long>(h) > (1ULL << 32)) die(path + ": image too large");
    f.get();  // the single whitespace after maxval
    im.w = static_cast<size_t>(w), im.h = static_cast<size_t>(h);
    im.px.resize(im.w * im.h * 3);
    f.read(reinterpret_cast<char*>(im.px.data()), static_cast<std::streamsize>(im.px.size()));
    if (static_cast<size_t>(f.gcount()) != im.px.size()) die(path + ": truncated pixel data");
    return im;
}

void savePpm(const std::string& path, const uint8_t* px, size_t w, size_t h) {
    std::ofstream f(path, std::ios::binary | std::ios::trunc);
    if (!f) die("cannot write " + path);
    f << "P6\n" << w << " " << h << "\n255\n";
    f.write(reinterpret_cast<const char*>(px), static_cast<std::streamsize>(w * h * 3));
    f.flush();
    if (!f) die("write failed: " + path);
}

std::vector<uint8_t> loadFile(const std::string& path) {
    std::error_code ec;
    if (!std::filesystem::is_regular_file(path, ec)) die("cannot open " + path);
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) die("cannot open " + path);
    const std::streamsize n = f.tellg();
    if (n < 0) die("cannot read " + path);
    std::vector<uint8_t> b(static_cast<size_t>(n));
    f.seekg(0);
    f.read(reinterpret_cast<char*>(b.data()), n);
    return b;
}

// `.mndict` (imageexperiments_b200/dictionary.py): "MNDICT1\0" | i32 n_base | i32 cols | i32 rows[n_base] | f32 base | f32 detail Y,U,V
mnenc_t* createContext(const std::string& exeDir) {
    const char* env = std::getenv("MNENC_DICT");
    // A `.mn` stream does not say which dictionary wrote it: decoding with another one "succeeds" and gives a wrong picture. So
    // whenever the dictionary is not the default one, say so once on stderr (the default is the export of the reference build).
    if (env)
        std::cerr << "Compression: note: dictionary " << (std::string(env) == "builtin" ? "computed by the library (MNENC_DICT=builtin, own codec revision)" : std::string("from ") + env)
                  << " - streams written with it decode only with the same dictionary, not with the default one or a reference build\n";
    if (env && std::string(env) == "builtin") {
        // the dictionary computed by the library itself (mnenc_build_dictionary): no reference build needed, but its KLT rows are
        // this library's own codec revision - encode and decode with the same setting
        std::vector<float> base(512 * 64), det[3];
        for (auto& d : det) d.resize(static_cast<size_t>(512) * 64 * 64);
        std::vector<int32_t> rows(512);
        int nBase = 0;
        if (mnenc_build_dictionary(kBlockSize, 0, base.data(), &nBase, rows.data(), det[0].data(), det[1].data(), det[2].data()) != MNENC_OK)
            dieLib("mnenc_build_dictionary");
        mnenc_t* ctx = nullptr;
        if (mnenc_create_auto(&ctx, kK, kBlockSize, base.data(), nBase, rows.data(), det[0].data(), det[1].data(), det[2].data()) != MNENC_OK)
            dieLib("mnenc_create_auto");
        return ctx;
    }
    const std::string path = env ? env : exeDir + "/data/dict_bs8.mndict";
    const std::vector<uint8_t> raw = loadFile(path);
    if (raw.size() < 16 || std::memcmp(raw.data(), "MNDICT1\0", 8) != 0) die(path + ": not a .mndict file");
    int32_t nBase, cols;
    std::memcpy(&nBase, raw.data() + 8, 4);
    std::memcpy(&cols, raw.data() + 12, 4);
    if (nBase <= 0 || cols != kBlockSize * kBlockSize) die(path + ": unexpected dictionary shape");
    size_t o = 16;
    const int32_t* rows = reinterpret_cast<const int32_t*>(raw.data() + o);
    o += 4 * static_cast<size_t>(nBase);
    if (o > raw.size()) die(path + ": truncated");
    size_t total = 0;
    for (int i = 0; i < nBase; ++i) total += static_cast<size_t>(rows[i]);
    const float* base = reinterpret_cast<const float*>(raw.data() + o);
    o += 4 * static_cast<size_t>(nBase) * cols;
    const float* det[3];
    for (auto& d : det) {
        d = reinterpret_cast<const float*>(raw.data() + o);
        o += 4 * total * cols;
    }
    if (o != raw.size()) die(path + ": size does not match its header");
    mnenc_t* ctx = nullptr;  // MNENC_DEVICE=n, or MNENC_DEVICES=0,1,... / all for one image over several GPUs
    if (mnenc_create_auto(&ctx, kK, kBlockSize, base, nBase, rows, det[0], det[1], det[2]) != MNENC_OK) dieLib("mnenc_create_auto");
    return ctx;
}

void setQuality(mnenc_t* ctx, const std::string& q) {
    std::vector<float> quant(3 * kK, 1.0f);  // "max": every quantiser 1 (Compression.cpp:104-114)
    if (q != "max" && mnenc_quant_tables(kK, kBlockSize, std::atof(q.c_str()), quant.data()) != MNENC_OK) dieLib("mnenc_quant_tables");
    if (mnenc_set_quant(ctx, quant.data()) != MNENC_OK) dieLib("mnenc_set_quant");
}

// compressed::calculatePSNR (CompressedImage.cpp:343-357): note the 3 * 255 peak.
double psnr(const Image& a, const uint8_t* b) {
    double se = 0.0;
    for (size_t x = 0; x < a.w; ++x)
        for (size_t y = 0; y < a.h; ++y) {
            const uint8_t* p = b + (y * a.w + x) * 3;
            const uint8_t* q = a.px.data() + (y * a.w + x) * 3;
            const double r = double(p[0]) - double(q[0]), g = double(p[1]) - double(q[1]), bl = double(p[2]) - double(q[2]);
            se += r * r + g * g + bl * bl;
        }
    const double mse = se / static_cast<double>(a.w * a.h);
    return 20.0 * std::log10(3.0 * 255.0) - 10.0 * std::log10(mse);
}

// Every regular file below `path`, sub-folders included (readFiles, Compression.cpp:31-52); sorted so that the report does
// not depend on the file system's directory order.
std::vector<std::string> readFiles(const std::string& path) {
    std::vector<std::string> results;
    std::error_code ec;
    for (std::filesystem::recursive_directory_iterator it(path, ec), end; !ec && it != end; it.increment(ec))
        if (it->is_regular_file(ec)) results.push_back(it->path().string());
    if (ec) die("cannot read folder " + path + ": " + ec.message());
    std::sort(results.begin(), results.end());
    return results;
}

void usage() {  // Compression.cpp:75-89, the modes kept
    std::cout << "To compress to file" << std::endl;
    std::cout << "usage: Compression -c <quality 1.0-8.0> <input PPM file path> <output compressed file path>" << std::endl;
    std::cout << "To decompress to file" << std::endl;
    std::cout << "usage: Compression -d <input compressed file path> <output PPM file path>" << std::endl;
    std::cout << "To display psnr and bpp of compression cycle" << std::endl;
    std::cout << "usage: Compression -n  <quality 1.0-8.0> <input PPM file path> [optional output PPM path]" << std::endl;
    std::cout << "Enumerate all .PPM files in folder and calclulate compression curves for them" << std::endl;
    std::cout << "Compression -g <input folder path> <outputs statistics file path>" << std::endl;
}

std::string dirOf(const char* argv0) {
    std::string s = argv0;
    const size_t k = s.find_last_of('/');
    return k == std::string::npos ? "." : s.substr(0, k);
}

}  // namespace

int main(int argc, char* argv[]) {
    if (argc < 2) {
        usage();
        return -1;
    }
    const std::string mode = argv[1], exeDir = dirOf(argv[0]);
    if (mode == "-c") {
        if (argc != 5) {
            usage();
            return -1;
        }
        std::cout << "create basis function" << std::endl;
        const Image in = loadPpm(argv[3]);
        mnenc_t* ctx = createContext(exeDir);
        setQuality(ctx, argv[2]);
        std::cout << std::format("start processing image {} using largest {} coefficients.", argv[3], kK) << std::endl;
        uint8_t* bytes = nullptr;
        size_t n = 0;
        if (mnenc_encode_rgb8(ctx, in.px.data(), in.w, in.h, 3 * in.w, &bytes, &n) != MNENC_OK) dieLib("mnenc_encode_rgb8");
        std::cout << std::format("bpp {} total KB {}", static_cast<double>(8ULL * n) / static_cast<double>(in.w * in.h), n / 1024ULL) << std::endl;
        std::ofstream out(argv[4], std::ios::out | std::ios::trunc | std::ios::binary);
        if (!out) die(std::string("cannot write ") + argv[4]);
        out.write(reinterpret_cast<const char*>(bytes), static_cast<std::streamsize>(n));
        out.flush();
        if (!out) die(std::string("write failed: ") + argv[4]);
        mnenc_free(bytes);
        mnenc_destroy(ctx);
    } else if (mode == "-d") {
        if (argc != 4) {
            usage();
            return -1;
        }
        const std::vector<uint8_t> buf = loadFile(argv[2]);
        mnenc_t* ctx = createContext(exeDir);
        uint8_t* px = nullptr;
        size_t w = 0, h = 0;
        if (mnenc_decode_rgb8(ctx, buf.data(), buf.size(), &px, &w, &h) != MNENC_OK) dieLib("mnenc_decode_rgb8");
        savePpm(argv[3], px, w, h);
        mnenc_free(px);
        mnenc_destroy(ctx);
    } else if (mode == "-n") {
        if (argc != 4 && argc != 5) {
            usage();
            return -1;
        }
        const auto start = std::chrono::system_clock::now();
        std::cout << "create basis function" << std::endl;
        const Image in = loadPpm(argv[3]);
        mnenc_t* ctx = createContext(exeDir);
        setQuality(ctx, argv[2]);
        std::cout << std::format("start processing image {} using largest {} coefficients.", argv[3], kK) << std::endl;
        uint8_t* bytes = nullptr;
        size_t n = 0;
        if (mnenc_encode_rgb8(ctx, in.px.data(), in.w, in.h, 3 * in.w, &bytes, &n) != MNENC_OK) dieLib("mnenc_encode_rgb8");
        uint8_t* px = nullptr;
        size_t w = 0, h = 0;
        if (mnenc_decode_rgb8(ctx, bytes, n, &px, &w, &h) != MNENC_OK) dieLib("mnenc_decode_rgb8");
        const auto end = std::chrono::system_clock::now();
        std::cout << std::chrono::duration_cast<std::chrono::milliseconds>(end - start) << std::endl;
        std::cout << std::format("PSNR {} bpp {} total KB {}", psnr(in, px), static_cast<double>(8ULL * n) / static_cast<double>(w * h), n / 1024ULL)
                  << std::endl;
        if (argc == 5) savePpm(argv[4], px, w, h);
        mnenc_free(px);
        mnenc_free(bytes);
        mnenc_destroy(ctx);
    } else if (mode == "-g") {
        if (argc != 4) {
            usage();
            return -1;
        }
        const std::vector<std::string> files = readFiles(argv[2]);
        std::ofstream fileOut(argv[3], std::ios::out | std::ios::trunc);
        if (!fileOut) die(std::string("cannot write ") + argv[3]);
        fileOut << "File, Mode, Quality, Size, BPP, PSNR" << std::endl;
        mnenc_t* ctx = nullptr;  // created at the first image: an empty folder needs no GPU
        for (size_t count = 0; count < files.size(); ++count) {
            const std::string& file = files[count];
            std::string lowerCaseName(file.size(), 0);
            std::transform(file.cbegin(), file.cend(), lowerCaseName.begin(), [](unsigned char c) { return static_cast<char>(std::tolower(c)); });
            if (!lowerCaseName.ends_with(".ppm")) continue;
            const Image in = loadPpm(file);
            if (!ctx) ctx = createContext(exeDir);
            // one context per quality in the reference (Compression.cpp:332); only the quantisers differ between them
            for (double bppAllocation = 8.0; bppAllocation >= 1.0; bppAllocation -= 1.0) {
                std::vector<float> quant(3 * kK);
                if (mnenc_quant_tables(kK, kBlockSize, bppAllocation, quant.data()) != MNENC_OK) dieLib("mnenc_quant_tables");
                if (mnenc_set_quant(ctx, quant.data()) != MNENC_OK) dieLib("mnenc_set_quant");
                std::cout << std::format("MP processing: {} {}%", file, (100L * static_cast<long>(count)) / static_cast<long>(files.size())) << std::endl;
                uint8_t* bytes = nullptr;
                size_t n = 0;
                if (mnenc_encode_rgb8(ctx, in.px.data(), in.w, in.h, 3 * in.w, &bytes, &n) != MNENC_OK) dieLib("mnenc_encode_rgb8");
                uint8_t* px = nullptr;
                size_t w = 0, h = 0;
                if (mnenc_decode_rgb8(ctx, bytes, n, &px, &w, &h) != MNENC_OK) dieLib("mnenc_decode_rgb8");
                const double p = psnr(in, px);
                mnenc_free(px);
                mnenc_free(bytes);
                const double bpp = static_cast<double>(8ULL * n) / static_cast<double>(in.w * in.h);
                std::cout << std::format("MP Quality {} PSNR {} bpp {} total KB {}", bppAllocation, p, bpp, n) << std::endl;  // bytes, as :353
                fileOut << std::format("{}, {}, {}, {}, {}, {}", file, "MP", bppAllocation, n, bpp, p) << std::endl;
            }
            fileOut.flush();
        }
        if (ctx) mnenc_destroy(ctx);
    } else {
        usage();
        return -1;
    }
    return 0;
}
