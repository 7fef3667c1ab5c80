// Caffe-free backing store for op::ArrayCpuGpu<T> -- compile INSTEAD of src/openpose/core/arrayCpuGpu.cpp.
// The reference's class is a facade over caffe::Blob<T> (include/openpose/core/arrayCpuGpu.hpp:14-106,
// src/openpose/core/arrayCpuGpu.cpp:26-891). Here the pimpl owns shape + pinned host buffer + device buffer and a
// lazy-sync head state with the observable behaviour of Caffe's SyncedMemory
// (3rdparty/caffe/src/caffe/syncedmem.cpp): cpu_data()/gpu_data() copy across only when the other side is newer,
// mutable_*() marks that side as the head, set_*_data() adopts an external (non-owned) pointer.
#include <openpose/core/arrayCpuGpu.hpp>
#include <cuda_runtime.h>
#include <cmath>
#include <numeric>
#include <openpose/utilities/errorAndLog.hpp>

namespace op
{
    template<typename T>
    struct ArrayCpuGpu<T>::ImplArrayCpuGpu
    {
        enum class Head { Uninitialized, AtCpu, AtGpu, Synced };
        std::vector<int> shape;
        size_t count = 0, capacity = 0;      // elements
        T* cpu = nullptr;
        T* gpu = nullptr;
        bool ownCpu = false, ownGpu = false;
        Head head = Head::Uninitialized;

        ~ImplArrayCpuGpu() { release(); }

        void release()
        {
            if (cpu && ownCpu)
                cudaFreeHost(cpu);
            if (gpu && ownGpu)
                cudaFree(gpu);
            cpu = nullptr;
            gpu = nullptr;
            ownCpu = ownGpu = false;
            capacity = 0;
            head = Head::Uninitialized;
        }
        void reshape(const std::vector<int>& newShape)
        {
            size_t n = 1;
            for (const auto d : newShape)
            {
                if (d < 0)
                    error("Negative dimension in ArrayCpuGpu::Reshape.", __LINE__, __FUNCTION__, __FILE__);
                n *= (size_t)d;
            }
            shape = newShape;
            count = n;
            if (n > capacity)      // like Blob::Reshape: memory only grows, contents are dropped when it does
            {
                const bool extCpu = cpu && !ownCpu, extGpu = gpu && !ownGpu;
                if (!extCpu && !extGpu)
                    release();
                capacity = n;
            }
        }
        void toCpu()
        {
            if (!cpu)
            {
                if (cudaMallocHost(reinterpret_cast<void**>(&cpu), sizeof(T) * std::max<size_t>(capacity, 1)) != cudaSuccess)
                    error("cudaMallocHost failed.", __LINE__, __FUNCTION__, __FILE__);
                ownCpu = true;
                if (head == Head::Uninitialized)
                    std::fill(cpu, cpu + capacity, T(0));
            }
            if (head == Head::AtGpu)
            {
                cudaMemcpy(cpu, gpu, sizeof(T) * count, cudaMemcpyDeviceToHost);
                head = Head::Synced;
            }
            else if (head == Head::Uninitialized)
                head = Head::AtCpu;
        }
        void toGpu()
        {
            if (!gpu)
            {
                if (cudaMalloc(reinterpret_cast<void**>(&gpu), sizeof(T) * std::max<size_t>(capacity, 1)) != cudaSuccess)
                    error("cudaMalloc failed.", __LINE__, __FUNCTION__, __FILE__);
                ownGpu = true;
                if (head == Head::Uninitialized)
                    cudaMemset(gpu, 0, sizeof(T) * capacity);
            }
            if (head == Head::AtCpu)
            {
                cudaMemcpy(gpu, cpu, sizeof(T) * count, cudaMemcpyHostToDevice);
                head = Head::Synced;
            }
            else if (head == Head::Uninitialized)
                head = Head::AtGpu;
        }
    };

    template<typename T>
    ArrayCpuGpu<T>::ArrayCpuGpu() : spImpl{new ImplArrayCpuGpu{}} {}

    template<typename T>
    ArrayCpuGpu<T>::ArrayCpuGpu(const void* caffeBlobTPtr) : spImpl{new ImplArrayCpuGpu{}}
    {
        // There is no Caffe blob on this path: a producer (op::NetB200) re-points a default-constructed
        // ArrayCpuGpu with Reshape + set_gpu_data instead.
        if (caffeBlobTPtr != nullptr)
            error("ArrayCpuGpu(const void* caffeBlobTPtr) needs Caffe; the B200 backend has no caffe::Blob.",
                  __LINE__, __FUNCTION__, __FILE__);
    }

    template<typename T>
    ArrayCpuGpu<T>::ArrayCpuGpu(const Array<T>& array, const bool copyFromGpu) : spImpl{new ImplArrayCpuGpu{}}
    {
        try
        {
            // reference: arrayCpuGpu.cpp:85-117 (4-D padding with leading ones, CPU copy only)
            if (copyFromGpu)
                error("Not implemented yet.", __LINE__, __FUNCTION__, __FILE__);
            std::vector<int> shape = array.getSize();
            while (shape.size() < 4)
                shape.insert(shape.begin(), 1);
            spImpl->reshape(shape);
            if (!array.empty())
                std::copy(array.getConstPtr(), array.getConstPtr() + array.getVolume(), mutable_cpu_data());
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    template<typename T>
    ArrayCpuGpu<T>::ArrayCpuGpu(const int num, const int channels, const int height, const int width) :
        spImpl{new ImplArrayCpuGpu{}}
    {
        spImpl->reshape({num, channels, height, width});
    }

    template<typename T>
    void ArrayCpuGpu<T>::Reshape(const int num, const int channels, const int height, const int width)
    {
        spImpl->reshape({num, channels, height, width});
    }
    template<typename T>
    void ArrayCpuGpu<T>::Reshape(const std::vector<int>& shape) { spImpl->reshape(shape); }

    template<typename T>
    std::string ArrayCpuGpu<T>::shape_string() const
    {
        std::string s;
        for (const auto d : spImpl->shape)
            s += std::to_string(d) + " ";
        return s + "(" + std::to_string(spImpl->count) + ")";
    }
    template<typename T>
    const std::vector<int>& ArrayCpuGpu<T>::shape() const { return spImpl->shape; }
    template<typename T>
    int ArrayCpuGpu<T>::shape(const int index) const { return spImpl->shape[CanonicalAxisIndex(index)]; }
    template<typename T>
    int ArrayCpuGpu<T>::num_axes() const { return (int)spImpl->shape.size(); }
    template<typename T>
    int ArrayCpuGpu<T>::count() const { return (int)spImpl->count; }
    template<typename T>
    int ArrayCpuGpu<T>::count(const int start_axis, const int end_axis) const
    {
        int c = 1;
        for (int i = start_axis; i < end_axis; ++i)
            c *= shape(i);
        return c;
    }
    template<typename T>
    int ArrayCpuGpu<T>::count(const int start_axis) const { return count(start_axis, num_axes()); }
    template<typename T>
    int ArrayCpuGpu<T>::CanonicalAxisIndex(const int axis_index) const
    {
        return axis_index < 0 ? axis_index + num_axes() : axis_index;
    }
    template<typename T>
    int ArrayCpuGpu<T>::LegacyShape(const int index) const
    {
        // Blob::LegacyShape: axes beyond num_axes() read as 1
        const int axes = num_axes();
        if (index >= axes || index < -axes)
            return 1;
        return shape(index);
    }
    template<typename T>
    int ArrayCpuGpu<T>::num() const { return LegacyShape(0); }
    template<typename T>
    int ArrayCpuGpu<T>::channels() const { return LegacyShape(1); }
    template<typename T>
    int ArrayCpuGpu<T>::height() const { return LegacyShape(2); }
    template<typename T>
    int ArrayCpuGpu<T>::width() const { return LegacyShape(3); }
    template<typename T>
    int ArrayCpuGpu<T>::offset(const int n, const int c, const int h, const int w) const
    {
        return ((n * channels() + c) * height() + h) * width() + w;
    }
    template<typename T>
    T ArrayCpuGpu<T>::data_at(const int n, const int c, const int h, const int w) const
    {
        return cpu_data()[offset(n, c, h, w)];
    }
    template<typename T>
    T ArrayCpuGpu<T>::diff_at(const int, const int, const int, const int) const { return T(0); }

    template<typename T>
    const T* ArrayCpuGpu<T>::cpu_data() const
    {
        spImpl->toCpu();
        return spImpl->cpu;
    }
    template<typename T>
    void ArrayCpuGpu<T>::set_cpu_data(T* data)
    {
        if (spImpl->cpu && spImpl->ownCpu)
            cudaFreeHost(spImpl->cpu);
        spImpl->cpu = data;
        spImpl->ownCpu = false;
        spImpl->head = ImplArrayCpuGpu::Head::AtCpu;
    }
    template<typename T>
    const int* ArrayCpuGpu<T>::gpu_shape() const
    {
        error("gpu_shape() is not available without Caffe.", __LINE__, __FUNCTION__, __FILE__);
        return nullptr;
    }
    template<typename T>
    const T* ArrayCpuGpu<T>::gpu_data() const
    {
        spImpl->toGpu();
        return spImpl->gpu;
    }
    template<typename T>
    void ArrayCpuGpu<T>::set_gpu_data(T* data)
    {
        if (spImpl->gpu && spImpl->ownGpu)
            cudaFree(spImpl->gpu);
        spImpl->gpu = data;
        spImpl->ownGpu = false;
        spImpl->head = ImplArrayCpuGpu::Head::AtGpu;
    }
    template<typename T>
    const T* ArrayCpuGpu<T>::cpu_diff() const { return nullptr; }   // inference only: no diffs
    template<typename T>
    const T* ArrayCpuGpu<T>::gpu_diff() const { return nullptr; }
    template<typename T>
    T* ArrayCpuGpu<T>::mutable_cpu_data()
    {
        spImpl->toCpu();
        spImpl->head = ImplArrayCpuGpu::Head::AtCpu;
        return spImpl->cpu;
    }
    template<typename T>
    T* ArrayCpuGpu<T>::mutable_gpu_data()
    {
        spImpl->toGpu();
        spImpl->head = ImplArrayCpuGpu::Head::AtGpu;
        return spImpl->gpu;
    }
    template<typename T>
    T* ArrayCpuGpu<T>::mutable_cpu_diff() { return nullptr; }
    template<typename T>
    T* ArrayCpuGpu<T>::mutable_gpu_diff() { return nullptr; }
    template<typename T>
    void ArrayCpuGpu<T>::Update() {}

    template<typename T>
    T ArrayCpuGpu<T>::asum_data() const
    {
        const T* p = cpu_data();
        T s = T(0);
        for (size_t i = 0; i < spImpl->count; ++i)
            s += (p[i] < T(0) ? T(0) - p[i] : p[i]);
        return s;
    }
    template<typename T>
    T ArrayCpuGpu<T>::asum_diff() const { return T(0); }
    template<typename T>
    T ArrayCpuGpu<T>::sumsq_data() const
    {
        const T* p = cpu_data();
        T s = T(0);
        for (size_t i = 0; i < spImpl->count; ++i)
            s += p[i] * p[i];
        return s;
    }
    template<typename T>
    T ArrayCpuGpu<T>::sumsq_diff() const { return T(0); }
    template<typename T>
    void ArrayCpuGpu<T>::scale_data(const T scale_factor)
    {
        T* p = mutable_cpu_data();
        for (size_t i = 0; i < spImpl->count; ++i)
            p[i] *= scale_factor;
    }
    template<typename T>
    void ArrayCpuGpu<T>::scale_diff(const T) {}

    COMPILE_TEMPLATE_BASIC_TYPES_CLASS(ArrayCpuGpu);
}
