// vulcan/buffer.h - owning device array (reference: include/vulcan/buffer.h:9-115,
// tests/buffer_test.cu:17-182): Resize grows the capacity exactly to the requested size and never
// shrinks it, Reserve re-allocates WITHOUT keeping the old contents, copies move GetSize() elements.
#pragma once

#include <cstddef>
#include <vulcan/device.h>

namespace vulcan
{

template <typename T>
class Buffer
{
  public:

    Buffer() : data_(nullptr), capacity_(0), size_(0) {}
    explicit Buffer(size_t size) : data_(nullptr), capacity_(0), size_(0) { Resize(size); }
    ~Buffer() { if (owned_) cudaFree(data_); }

    const T* GetData() const { return data_; }
    T* GetData() { return data_; }
    size_t GetSize() const { return size_; }
    size_t GetCapacity() const { return capacity_; }
    size_t IsEmpty() const { return size_ == 0; }

    void Resize(size_t size)
    {
      if (size > capacity_) Reserve(size);
      size_ = size;
    }

    void Reserve(size_t capacity)
    {
      if (capacity > capacity_)
      {
        VULCAN_ASSERT_MSG(owned_, "cannot grow a buffer that views library memory");
        CUDA_DEBUG(cudaFree(data_));
        data_ = nullptr;
        CUDA_ASSERT(cudaMalloc((void**)&data_, sizeof(T) * capacity));
        capacity_ = capacity;
      }
    }

    void CopyFromDevice(const T* buffer) { Copy(data_, buffer, cudaMemcpyDeviceToDevice); }
    void CopyToDevice(T* buffer) const { Copy(buffer, data_, cudaMemcpyDeviceToDevice); }
    void CopyFromHost(const T* buffer) { Copy(data_, buffer, cudaMemcpyHostToDevice); }
    void CopyToHost(T* buffer) const { Copy(buffer, data_, cudaMemcpyDeviceToHost); }

    // Not in the reference: the buffer becomes a non-owning view of memory that the C-ABI
    // library owns (vulcan::Volume exposes its pools this way, see volume.h).
    void Adopt(T* data, size_t capacity, size_t size)
    {
      if (owned_) cudaFree(data_);
      owned_ = false;
      data_ = data; capacity_ = capacity; size_ = size;
    }

    // Not in the reference: logical size inside the current capacity (visible-block count)
    void SetSize(size_t size) { VULCAN_ASSERT(size <= capacity_); size_ = size; }

  private:

    Buffer(const Buffer&);
    Buffer& operator=(const Buffer&);

    void Copy(void* dst, const void* src, cudaMemcpyKind kind) const
    {
      if (size_ == 0) return;
      CUDA_ASSERT(cudaMemcpy(dst, src, sizeof(T) * size_, kind));
    }

  protected:

    T* data_;
    size_t capacity_;
    size_t size_;
    bool owned_ = true;
};

} // namespace vulcan
