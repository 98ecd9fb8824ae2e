// vulcan/image.h - owning device images (reference: include/vulcan/image.h:11-273): Image = float,
// ColorImage = packed 3 x float per pixel, row-major y*W + x; Resize re-allocates only when the
// pixel count changes.  The reference's Load/Save go through OpenCV (absent here and out of scope);
// this header offers raw host copies and portable float maps (PFM) instead.
#pragma once

#include <cstdio>
#include <string>
#include <vector>
#include <vulcan/device.h>
#include <vulcan/matrix.h>

namespace vulcan
{

namespace detail
{

template <typename T, int CHANNELS>
class DeviceImage
{
  public:

    DeviceImage() : size_(0, 0), data_(nullptr) {}
    DeviceImage(int w, int h) : size_(0, 0), data_(nullptr) { Resize(w, h); }
    ~DeviceImage() { cudaFree(data_); }

    int GetWidth() const { return size_[0]; }
    int GetHeight() const { return size_[1]; }
    const Vector2i& GetSize() const { return size_; }
    int GetTotal() const { return size_[0] * size_[1]; }
    int GetBytes() const { return (int)sizeof(T) * GetTotal(); }
    const T* GetData() const { return data_; }
    T* GetData() { return data_; }

    void Resize(int w, int h) { Resize(Vector2i(w, h)); }

    void Resize(const Vector2i& size)
    {
      VULCAN_DEBUG(size[0] >= 0 && size[1] >= 0);
      const int old_total = GetTotal();
      size_ = size;
      if (GetTotal() != old_total)
      {
        CUDA_DEBUG(cudaFree(data_));
        data_ = nullptr;
        if (GetTotal() > 0) CUDA_ASSERT(cudaMalloc((void**)&data_, GetBytes()));
      }
    }

    // raw host <-> device copies of the whole image (not in the reference, which loads via OpenCV)
    void CopyFromHost(const void* host) { if (GetTotal()) CUDA_ASSERT(cudaMemcpy(data_, host, GetBytes(), cudaMemcpyHostToDevice)); }
    void CopyToHost(void* host) const { if (GetTotal()) CUDA_ASSERT(cudaMemcpy(host, data_, GetBytes(), cudaMemcpyDeviceToHost)); }

    // portable float map, bottom-up rows as the format demands ("Pf" grey / "PF" colour)
    void Save(const std::string& file) const
    {
      std::vector<float> host((size_t)GetTotal() * CHANNELS);
      CopyToHost(host.data());
      FILE* f = fopen(file.c_str(), "wb");
      VULCAN_ASSERT_MSG(f != nullptr, "unable to save file");
      fprintf(f, "%s\n%d %d\n-1.0\n", CHANNELS == 3 ? "PF" : "Pf", GetWidth(), GetHeight());
      for (int y = GetHeight() - 1; y >= 0; --y)
        fwrite(host.data() + (size_t)y * GetWidth() * CHANNELS, sizeof(float), (size_t)GetWidth() * CHANNELS, f);
      fclose(f);
    }

    void Load(const std::string& file, float scale = 1)
    {
      FILE* f = fopen(file.c_str(), "rb");
      VULCAN_ASSERT_MSG(f != nullptr, "unable to load file");
      char magic[3] = {0, 0, 0};
      int w = 0, h = 0;
      float endian = 0;
      const bool ok = fscanf(f, "%2s %d %d %f", magic, &w, &h, &endian) == 4 && fgetc(f) != EOF &&
                      std::string(magic) == (CHANNELS == 3 ? "PF" : "Pf") && endian < 0;
      if (!ok) { fclose(f); VULCAN_THROW("unable to load file"); }
      std::vector<float> host((size_t)w * h * CHANNELS);
      for (int y = h - 1; y >= 0; --y)
        if (fread(host.data() + (size_t)y * w * CHANNELS, sizeof(float), (size_t)w * CHANNELS, f) != (size_t)w * CHANNELS)
        { fclose(f); VULCAN_THROW("unable to load file"); }
      fclose(f);
      for (float& v : host) v *= scale;
      Resize(w, h);
      CopyFromHost(host.data());
    }

  private:

    DeviceImage(const DeviceImage&);
    DeviceImage& operator=(const DeviceImage&);

  protected:

    Vector2i size_;
    T* data_;
};

} // namespace detail

class Image : public detail::DeviceImage<float, 1>
{
  public:
    Image() {}
    Image(int w, int h) : detail::DeviceImage<float, 1>(w, h) {}
    // half-size image, nearest sample or 2x2 average (src/image.cu:183-212)
    void Downsample(Image& image, bool nearest) const;
};

class ColorImage : public detail::DeviceImage<Vector3f, 3>
{
  public:
    ColorImage() {}
    ColorImage(int w, int h) : detail::DeviceImage<Vector3f, 3>(w, h) {}
    // src/image.cu:234-263
    void Downsample(ColorImage& image, bool nearest) const;
};

} // namespace vulcan
