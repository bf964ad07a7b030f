// rgbd_file_b200_provider.hpp -- SURVEY 8(f) row f3: the input side of the hot path.
//
//   mslam::RgbdFileB200Provider  replaces  mslam::RgbdFileProvider
//       (include/modular_slam/sensors/rgbd_file_provider.hpp:23-48, src/lib/modular_slam/rgbd_file_provider.cpp:42-107)
//   mslam::readTumRgbdDataset / tumRgbdCameraParams / getImages   (rgbd_file_provider.cpp:109-145, :26-40)
//
// Same interface (DataProviderInterface<RgbdFrame>: init / fetch / recentData) and the same results: RgbFrame::data is
// what cv::imread(path) returns (3 interleaved bytes per pixel in B, G, R order, rgbd_file_provider.cpp:66,94) and
// DepthFrame::data what cv::imread(path, IMREAD_ANYDEPTH) returns (one u16 per pixel, :67,95).  OpenCV's imgcodecs is
// not in this image, so PNG decoding is done here (zlib inflate + the five PNG row filters); JPEG is not supported
// (the reference's datasets are PNG; a .jpg path makes fetch() fail the way an unreadable file does).
//
// What it adds for the GPU path: fetchBatch() decodes the next n frames on a pool of host threads straight into
// PINNED host memory (mslam_b200_host_alloc), so that the detector's host entry point (mslam_b200_detect_batch: chunked
// copy-in / kernels / copy-out pipeline) streams them to the GPU at full PCIe rate without a staging copy.
#pragma once
#ifdef MSLAM_B200_WITH_REFERENCE_HEADERS
#include "modular_slam/sensors/rgbd_file_provider.hpp"
#else
#include "mslam_shim.hpp"
#endif
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace mslam
{

#ifndef MSLAM_B200_WITH_REFERENCE_HEADERS
struct RgbdFilePaths // rgbd_file_provider.hpp:16-21
{
    std::vector<std::string> rgbPaths;
    std::vector<std::string> depthPaths;
    std::vector<double> timestamps;
};
#endif

// PNG -> pixels, with cv::imread's conversions.  Returns false on anything unreadable (missing file, not a PNG, bad CRC-less
// stream, Adam7 interlace).  `out` must hold width * height * 3 bytes (bgr) / width * height u16 (depth).
bool readPngInfo(const std::string& path, int* width, int* height, int* bitDepth, int* colorType);
// cv::imread(path): 8-bit, 3 channels, B G R (16-bit samples keep their high byte, alpha is dropped, gray / palette expand)
bool readPngBgr8(const std::string& path, std::uint8_t* out, int width, int height);
// cv::imread(path, cv::IMREAD_ANYDEPTH): single channel, 16-bit files keep 16 bits, 8-bit files widen without scaling
// is NOT what OpenCV does (it returns CV_8U) -- the reference then reinterprets the buffer as CV_16UC1
// (rgbd_file_provider.cpp:92), which only works for 16-bit depth maps; this reader refuses anything but 16-bit gray.
bool readPngGray16(const std::string& path, std::uint16_t* out, int width, int height);

class RgbdFileB200Provider : public DataProviderInterface<RgbdFrame>
{
  public:
    RgbdFileB200Provider(RgbdFilePaths paths, const CameraParameters& cameraParams);
    ~RgbdFileB200Provider() override;
    RgbdFileB200Provider() = delete;
    RgbdFileB200Provider(const RgbdFileB200Provider&) = delete;
    RgbdFileB200Provider& operator=(const RgbdFileB200Provider&) = delete;

    bool init() override;
    bool fetch() override;
    std::shared_ptr<RgbdFrame> recentData() const override;

    // The batched path.  Decodes up to maxFrames following frames (all of the first one's size) into pinned host memory
    // owned by the provider and valid until the next fetchBatch(); returns how many were read (0 at the end of the
    // sequence or when the next frame is unreadable).  rgb[i] / depth[i] point at frame i.
    struct Batch
    {
        int count = 0;
        int width = 0, height = 0;
        std::vector<const std::uint8_t*> rgb;
        std::vector<const std::uint16_t*> depth;
        std::vector<double> timestamps;
    };
    Batch fetchBatch(int maxFrames, int threads = 0);
    std::size_t position() const { return currentIndex; }

  private:
    RgbdFilePaths filePaths;
    std::size_t currentIndex = 0;
    std::shared_ptr<RgbdFrame> recentFrame;
    std::shared_ptr<RgbdFrame> decode(std::size_t index) const;
    CameraParameters cameraParameters;
    void* pinned = nullptr; // mslam_b200_host_alloc
    std::size_t pinnedBytes = 0;
};

RgbdFilePaths readTumRgbdDataset(const std::string& tumFile); // rgbd_file_provider.cpp:109-134
CameraParameters tumRgbdCameraParams();                       // rgbd_file_provider.cpp:136-145
std::vector<std::string> getImages(const std::string& rootDir); // rgbd_file_provider.cpp:26-40

} // namespace mslam
