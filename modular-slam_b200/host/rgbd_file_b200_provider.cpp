// rgbd_file_b200_provider.cpp -- see rgbd_file_b200_provider.hpp.
#include "rgbd_file_b200_provider.hpp"

#include <dirent.h>
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <thread>

#include "../../include/mslam_b200.h"

namespace mslam
{

// ---- PNG (ISO/IEC 15948): signature, IHDR, IDAT* (one zlib stream), IEND; five row filters; no Adam7 ----------------
namespace
{
struct PngImage
{
    int width = 0, height = 0, bitDepth = 0, colorType = 0, interlace = 0;
    std::vector<std::uint8_t> palette;   // PLTE, 3 bytes per entry
    std::vector<std::uint8_t> idat;      // concatenated IDAT payloads
    std::vector<std::uint8_t> pixels;    // unfiltered scanlines, `stride` bytes per row
    std::size_t stride = 0;
    int channels = 0;
};

inline std::uint32_t be32(const std::uint8_t* p)
{
    return (std::uint32_t)p[0] << 24 | (std::uint32_t)p[1] << 16 | (std::uint32_t)p[2] << 8 | (std::uint32_t)p[3];
}

bool readFile(const std::string& path, std::vector<std::uint8_t>* out)
{
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if(!f)
        return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if(n <= 0)
    {
        std::fclose(f);
        return false;
    }
    out->resize((std::size_t)n);
    const std::size_t got = std::fread(out->data(), 1, (std::size_t)n, f);
    std::fclose(f);
    return got == (std::size_t)n;
}

bool parseChunks(const std::vector<std::uint8_t>& file, PngImage* img, bool headerOnly)
{
    static const std::uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if(file.size() < 8 + 25 || std::memcmp(file.data(), sig, 8) != 0)
        return false;
    std::size_t pos = 8;
    bool haveHeader = false;
    while(pos + 12 <= file.size())
    {
        const std::uint32_t len = be32(&file[pos]);
        const std::uint8_t* type = &file[pos + 4];
        if(len > file.size() - pos - 12)
            return false;
        const std::uint8_t* data = &file[pos + 8];
        if(!std::memcmp(type, "IHDR", 4))
        {
            if(len != 13)
                return false;
            img->width = (int)be32(data);
            img->height = (int)be32(data + 4);
            img->bitDepth = data[8];
            img->colorType = data[9];
            img->interlace = data[12];
            if(img->width <= 0 || img->height <= 0 || data[10] != 0 || data[11] != 0)
                return false;
            haveHeader = true;
            if(headerOnly)
                return true;
        }
        else if(!std::memcmp(type, "PLTE", 4))
            img->palette.assign(data, data + len);
        else if(!std::memcmp(type, "IDAT", 4))
            img->idat.insert(img->idat.end(), data, data + len);
        else if(!std::memcmp(type, "IEND", 4))
            break;
        pos += 12 + (std::size_t)len;
    }
    return haveHeader && !img->idat.empty();
}

inline int paeth(int a, int b, int c)
{
    const int p = a + b - c;
    const int pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

bool decodePng(const std::string& path, PngImage* img)
{
    std::vector<std::uint8_t> file;
    if(!readFile(path, &file) || !parseChunks(file, img, false))
        return false;
    if(img->interlace != 0)
        return false; // Adam7: not produced by the datasets this provider reads
    switch(img->colorType)
    {
    case 0: img->channels = 1; break;
    case 2: img->channels = 3; break;
    case 3: img->channels = 1; break;
    case 4: img->channels = 2; break;
    case 6: img->channels = 4; break;
    default: return false;
    }
    const int bd = img->bitDepth;
    if(!(bd == 8 || bd == 16 || ((img->colorType == 0 || img->colorType == 3) && (bd == 1 || bd == 2 || bd == 4))))
        return false;
    if(img->colorType == 3 && (bd == 16 || img->palette.size() < 3))
        return false;
    const std::size_t bitsPerPixel = (std::size_t)img->channels * bd;
    img->stride = ((std::size_t)img->width * bitsPerPixel + 7) / 8;
    const std::size_t bpp = std::max<std::size_t>(1, bitsPerPixel / 8); // filter distance in bytes
    std::vector<std::uint8_t> raw((img->stride + 1) * (std::size_t)img->height);
    uLongf rawLen = (uLongf)raw.size();
    if(uncompress(raw.data(), &rawLen, img->idat.data(), (uLong)img->idat.size()) != Z_OK || rawLen != raw.size())
        return false;
    img->pixels.resize(img->stride * (std::size_t)img->height);
    const std::size_t n = img->stride;
    for(int y = 0; y < img->height; ++y)
    {
        const std::uint8_t* in = &raw[(n + 1) * (std::size_t)y];
        std::uint8_t* cur = &img->pixels[n * (std::size_t)y];
        const std::uint8_t* up = y ? cur - n : nullptr;
        const int filter = in[0];
        ++in;
        switch(filter)
        {
        case 0: std::memcpy(cur, in, n); break;
        case 1:
            for(std::size_t i = 0; i < n; ++i)
                cur[i] = (std::uint8_t)(in[i] + (i >= bpp ? cur[i - bpp] : 0));
            break;
        case 2:
            for(std::size_t i = 0; i < n; ++i)
                cur[i] = (std::uint8_t)(in[i] + (up ? up[i] : 0));
            break;
        case 3:
            for(std::size_t i = 0; i < n; ++i)
                cur[i] = (std::uint8_t)(in[i] + (((i >= bpp ? cur[i - bpp] : 0) + (up ? up[i] : 0)) >> 1));
            break;
        case 4:
            for(std::size_t i = 0; i < n; ++i)
                cur[i] = (std::uint8_t)(in[i] + paeth(i >= bpp ? cur[i - bpp] : 0, up ? up[i] : 0, (up && i >= bpp) ? up[i - bpp] : 0));
            break;
        default: return false;
        }
    }
    return true;
}

// sample `c` of pixel x in an unfiltered row, as an 8-bit value the way libpng hands it to cv::imread's 8-bit path:
// 16-bit samples are stripped to their high byte (png_set_strip_16), sub-byte gray is scaled up (png_set_expand_gray_1_2_4_to_8)
inline int sample8(const PngImage& img, const std::uint8_t* row, int x, int c)
{
    const int bd = img.bitDepth;
    if(bd == 8)
        return row[(std::size_t)x * img.channels + c];
    if(bd == 16)
        return row[((std::size_t)x * img.channels + c) * 2];
    const int perByte = 8 / bd;
    const int v = (row[x / perByte] >> (8 - bd * (x % perByte + 1))) & ((1 << bd) - 1);
    return img.colorType == 3 ? v : v * 255 / ((1 << bd) - 1);
}
} // namespace

bool readPngInfo(const std::string& path, int* width, int* height, int* bitDepth, int* colorType)
{
    std::vector<std::uint8_t> head(64);
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if(!f)
        return false;
    head.resize(std::fread(head.data(), 1, head.size(), f));
    std::fclose(f);
    PngImage img;
    if(!parseChunks(head, &img, true))
        return false;
    if(width) *width = img.width;
    if(height) *height = img.height;
    if(bitDepth) *bitDepth = img.bitDepth;
    if(colorType) *colorType = img.colorType;
    return true;
}

bool readPngBgr8(const std::string& path, std::uint8_t* out, int width, int height)
{
    PngImage img;
    if(!decodePng(path, &img) || img.width != width || img.height != height)
        return false;
    for(int y = 0; y < height; ++y)
    {
        const std::uint8_t* row = &img.pixels[img.stride * (std::size_t)y];
        std::uint8_t* o = out + (std::size_t)y * width * 3;
        for(int x = 0; x < width; ++x, o += 3)
        {
            int r, g, b;
            if(img.colorType == 3)
            {
                const std::size_t idx = (std::size_t)sample8(img, row, x, 0) * 3;
                if(idx + 2 < img.palette.size())
                    r = img.palette[idx], g = img.palette[idx + 1], b = img.palette[idx + 2];
                else
                    r = g = b = 0;
            }
            else if(img.channels <= 2)
                r = g = b = sample8(img, row, x, 0);
            else
                r = sample8(img, row, x, 0), g = sample8(img, row, x, 1), b = sample8(img, row, x, 2);
            o[0] = (std::uint8_t)b, o[1] = (std::uint8_t)g, o[2] = (std::uint8_t)r; // cv::imread: B G R
        }
    }
    return true;
}

bool readPngGray16(const std::string& path, std::uint16_t* out, int width, int height)
{
    PngImage img;
    if(!decodePng(path, &img) || img.width != width || img.height != height || img.bitDepth != 16 || img.colorType != 0)
        return false;
    for(int y = 0; y < height; ++y)
    {
        const std::uint8_t* row = &img.pixels[img.stride * (std::size_t)y];
        std::uint16_t* o = out + (std::size_t)y * width;
        for(int x = 0; x < width; ++x)
            o[x] = (std::uint16_t)(row[2 * x] << 8 | row[2 * x + 1]); // PNG samples are big-endian
    }
    return true;
}

// ---- dataset listing (rgbd_file_provider.cpp:16-40, :109-145) ---------------------------------------------------------
static bool hasImageExtension(const std::string& path)
{
    const std::size_t dot = path.rfind('.');
    if(dot == std::string::npos || path.find('/', dot) != std::string::npos)
        return false;
    const std::string ext = path.substr(dot);
    return ext == ".png" || ext == ".jpg" || ext == ".jpeg";
}

std::vector<std::string> getImages(const std::string& rootDir)
{
    std::vector<std::string> imagePaths;
    if(DIR* d = opendir(rootDir.c_str()))
    {
        while(dirent* e = readdir(d))
        {
            const std::string name = e->d_name;
            if(name != "." && name != ".." && hasImageExtension(name))
                imagePaths.push_back(rootDir + (rootDir.empty() || rootDir.back() == '/' ? "" : "/") + name);
        }
        closedir(d);
    }
    std::sort(imagePaths.begin(), imagePaths.end());
    return imagePaths;
}

// One association line is "<rgb time> <rgb file> <depth time> <depth file>" (associate.py of the TUM RGB-D tools); the
// reference keeps the first time stamp and the two file names, relative to the list's directory, and skips lines that do
// not have all four fields (rgbd_file_provider.cpp:109-134).  Tokenised by hand here: no stream state to reason about.
namespace
{
bool nextToken(const std::string& text, std::size_t* at, std::string* token)
{
    const char* blanks = " \t\r\n";
    const std::size_t first = text.find_first_not_of(blanks, *at);
    if(first == std::string::npos)
        return false;
    const std::size_t last = text.find_first_of(blanks, first);
    *token = text.substr(first, last == std::string::npos ? std::string::npos : last - first);
    *at = last == std::string::npos ? text.size() : last;
    return true;
}

bool toDouble(const std::string& token, double* value)
{
    char* end = nullptr;
    *value = std::strtod(token.c_str(), &end);
    return end != token.c_str(); // operator>> semantics: a leading number is enough ("1.5x" reads 1.5)
}
} // namespace

RgbdFilePaths readTumRgbdDataset(const std::string& tumFile)
{
    RgbdFilePaths result;
    std::string directory;
    if(const std::size_t cut = tumFile.find_last_of('/'); cut != std::string::npos)
        directory = tumFile.substr(0, cut);
    // what std::filesystem::path::operator/ does for the two cases that occur: an absolute name stands alone
    const auto resolve = [&directory](const std::string& name) {
        return (name.front() == '/' || directory.empty()) ? name : directory + '/' + name;
    };
    std::ifstream list(tumFile);
    for(std::string row; std::getline(list, row);)
    {
        std::size_t cursor = 0;
        std::string field[4];
        bool complete = true;
        for(std::string& f : field)
            complete = complete && nextToken(row, &cursor, &f);
        double stamp = 0.0;
        if(!complete || !toDouble(field[0], &stamp))
            continue;
        result.timestamps.push_back(stamp);
        result.rgbPaths.push_back(resolve(field[1]));
        result.depthPaths.push_back(resolve(field[3]));
    }
    return result;
}

CameraParameters tumRgbdCameraParams()
{
    CameraParameters cameraParams{};
    cameraParams.focal[0] = 525, cameraParams.focal[1] = 525;
    cameraParams.principalPoint[0] = 319.5f, cameraParams.principalPoint[1] = 239.5f;
    cameraParams.factor = 1.f / 5000.f;
    return cameraParams;
}

// ---- the provider -----------------------------------------------------------------------------------------------------
RgbdFileB200Provider::RgbdFileB200Provider(RgbdFilePaths paths, const CameraParameters& params)
    : filePaths(std::move(paths)), cameraParameters(params)
{
}

RgbdFileB200Provider::~RgbdFileB200Provider()
{
    if(pinned)
        mslam_b200_host_free(pinned);
}

bool RgbdFileB200Provider::init()
{
    // a usable sequence has at least one frame and a depth image for every colour image (rgbd_file_provider.cpp:50-55)
    const std::size_t colour = filePaths.rgbPaths.size();
    return colour != 0 && colour == filePaths.depthPaths.size();
}

// decode frame `index` of the sequence into a fresh RgbdFrame; null when it does not exist or either image is unreadable
std::shared_ptr<RgbdFrame> RgbdFileB200Provider::decode(std::size_t index) const
{
    if(index >= filePaths.rgbPaths.size() || index >= filePaths.depthPaths.size())
        return nullptr;
    const std::string &colourFile = filePaths.rgbPaths[index], &depthFile = filePaths.depthPaths[index];
    Size colour{0, 0}, depth{0, 0};
    if(!readPngInfo(colourFile, &colour.width, &colour.height, nullptr, nullptr) ||
       !readPngInfo(depthFile, &depth.width, &depth.height, nullptr, nullptr))
        return nullptr;
    auto out = std::make_shared<RgbdFrame>();
    out->rgb.size = colour;
    out->rgb.data.resize((std::size_t)colour.width * colour.height * 3);
    out->depth.size = depth;
    out->depth.data.resize((std::size_t)depth.width * depth.height);
    out->depth.cameraParameters = cameraParameters;
    out->timestamp.timePoint = index < filePaths.timestamps.size() ? filePaths.timestamps[index] : 0.0;
    const bool readable = readPngBgr8(colourFile, out->rgb.data.data(), colour.width, colour.height) &&
                          readPngGray16(depthFile, out->depth.data.data(), depth.width, depth.height);
    return readable ? out : nullptr;
}

// The reference's contract (rgbd_file_provider.cpp:57-107): true and a new recentData() while frames remain; past the end,
// or on an image that cannot be read, recentData() becomes null and the position does not advance.
bool RgbdFileB200Provider::fetch()
{
    recentFrame = decode(currentIndex);
    if(recentFrame)
        ++currentIndex;
    return recentFrame != nullptr;
}

std::shared_ptr<RgbdFrame> RgbdFileB200Provider::recentData() const
{
    return recentFrame;
}

RgbdFileB200Provider::Batch RgbdFileB200Provider::fetchBatch(int maxFrames, int threads)
{
    Batch b;
    const std::size_t total = std::min(filePaths.rgbPaths.size(), filePaths.depthPaths.size());
    if(maxFrames < 1 || currentIndex >= total)
        return b;
    int w = 0, h = 0;
    if(!readPngInfo(filePaths.rgbPaths[currentIndex], &w, &h, nullptr, nullptr))
        return b;
    const int want = (int)std::min<std::size_t>((std::size_t)maxFrames, total - currentIndex);
    const std::size_t rgbBytes = (std::size_t)w * h * 3, depthBytes = (std::size_t)w * h * 2;
    const std::size_t frameBytes = (rgbBytes + depthBytes + 255) / 256 * 256;
    if(pinnedBytes < frameBytes * (std::size_t)want)
    {
        if(pinned)
            mslam_b200_host_free(pinned);
        pinned = nullptr;
        pinnedBytes = 0;
        if(mslam_b200_host_alloc(&pinned, frameBytes * (std::size_t)want) != MSLAM_B200_OK)
            return b; // no CUDA runtime: the batched path has no pageable fallback
        pinnedBytes = frameBytes * (std::size_t)want;
    }
    std::vector<char> ok((std::size_t)want, 0);
    std::atomic<int> next{0};
    const std::size_t first = currentIndex;
    auto work = [&]() {
        for(int i = next.fetch_add(1); i < want; i = next.fetch_add(1))
        {
            std::uint8_t* base = static_cast<std::uint8_t*>(pinned) + frameBytes * (std::size_t)i;
            ok[(std::size_t)i] = readPngBgr8(filePaths.rgbPaths[first + (std::size_t)i], base, w, h) &&
                                 readPngGray16(filePaths.depthPaths[first + (std::size_t)i],
                                               reinterpret_cast<std::uint16_t*>(base + rgbBytes), w, h);
        }
    };
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    nt = std::max(1, std::min(nt, want));
    std::vector<std::thread> pool;
    for(int t = 1; t < nt; ++t)
        pool.emplace_back(work);
    work();
    for(auto& t : pool)
        t.join();
    int n = 0;
    while(n < want && ok[(std::size_t)n]) // stop at the first unreadable frame, like fetch()
        ++n;
    b.count = n;
    b.width = w, b.height = h;
    for(int i = 0; i < n; ++i)
    {
        const std::uint8_t* base = static_cast<const std::uint8_t*>(pinned) + frameBytes * (std::size_t)i;
        b.rgb.push_back(base);
        b.depth.push_back(reinterpret_cast<const std::uint16_t*>(base + rgbBytes));
        b.timestamps.push_back(first + (std::size_t)i < filePaths.timestamps.size() ? filePaths.timestamps[first + (std::size_t)i] : 0.0);
    }
    currentIndex += (std::size_t)n;
    return b;
}

} // namespace mslam
