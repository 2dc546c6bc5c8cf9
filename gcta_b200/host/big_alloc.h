// big_alloc.h - host buffers of gigabytes (the packed genotypes, the GRM triangles): anonymous mappings with
// MADV_HUGEPAGE, so that where transparent huge pages are in "madvise" mode the first touch of the buffer - by the
// reader's threads, by the library's copy-out threads - faults per 2 MB instead of per 4 KB.  Measured in the build
// container (8 threads touching 8 GiB for the first time): 2.4 s -> 0.6 s.  Never zero-filled by us; small requests go
// to operator new.
#ifndef GCTA_B200_BIG_ALLOC_H
#define GCTA_B200_BIG_ALLOC_H

#include <sys/mman.h>

#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <new>
#include <utility>

namespace big_alloc {

constexpr size_t THRESHOLD = 64u << 20;
// GCTA_B200_BIG_ALLOC=0: everything through operator new (to compare)
inline bool enabled()
{
    static const bool on = !(getenv("GCTA_B200_BIG_ALLOC") && atoi(getenv("GCTA_B200_BIG_ALLOC")) == 0);
    return on;
}

inline void* get(size_t bytes)
{
    if (bytes < THRESHOLD || !enabled()) return ::operator new(bytes);
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) throw std::bad_alloc();
#ifdef MADV_HUGEPAGE
    madvise(p, bytes, MADV_HUGEPAGE);   // a hint; failure (no THP) changes nothing
#endif
    return p;
}
inline void release(void* p, size_t bytes)
{
    if (!p) return;
    if (bytes < THRESHOLD || !enabled()) ::operator delete(p);
    else munmap(p, bytes);
}

// an uninitialised array of n trivially constructible elements
template <typename T>
struct Array {
    T* p = nullptr;
    size_t n = 0;
    Array() = default;
    explicit Array(size_t count) : p(count ? static_cast<T*>(big_alloc::get(count * sizeof(T))) : nullptr), n(count) {}
    Array(const Array&) = delete;
    Array& operator=(const Array&) = delete;
    Array(Array&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    Array& operator=(Array&& o) noexcept { if (this != &o) { reset(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; } return *this; }
    ~Array() { reset(); }
    void reset() { big_alloc::release(p, n * sizeof(T)); p = nullptr; n = 0; }
    T* get() const { return p; }
};

// std allocator over the same, leaving trivially constructible elements uninitialised (std::vector::resize does not
// write the bytes the reader is about to fill)
template <typename T>
struct allocator {
    typedef T value_type;
    allocator() = default;
    template <typename U> allocator(const allocator<U>&) noexcept {}
    template <typename U> struct rebind { typedef allocator<U> other; };
    T* allocate(size_t count) { return static_cast<T*>(big_alloc::get(count * sizeof(T))); }
    void deallocate(T* p, size_t count) noexcept { big_alloc::release(p, count * sizeof(T)); }
    template <typename U> void construct(U* p) noexcept { ::new (static_cast<void*>(p)) U; }
    template <typename U, typename... Args> void construct(U* p, Args&&... args) { ::new (static_cast<void*>(p)) U(std::forward<Args>(args)...); }
    template <typename U> bool operator==(const allocator<U>&) const noexcept { return true; }
    template <typename U> bool operator!=(const allocator<U>&) const noexcept { return false; }
};

}  // namespace big_alloc

#endif
