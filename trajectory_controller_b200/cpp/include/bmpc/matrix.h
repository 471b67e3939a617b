// bmpc::matrix<T,NR,NC> — the small fixed-size, row-major, by-value matrix the dlib::mpc-compatible facade
// (bmpc/mpc.h) takes and returns.  It offers exactly the slice of dlib::matrix the reference module and dlib's
// own MPC test use on the hot path:
//   storage        row-major, on the stack, no heap      (dlib_files/dlib/matrix/matrix_data_layout.h:131-155)
//   m(r,c), m(i)   element access                        (dlib_files/dlib/matrix/matrix.h:1063-1135)
//   m = a, b, c;   comma initialiser; `m = s` alone fills every element with s (matrix.h:1500-1620)
//   + - * (matrix, scalar)   plain per-element / dot-product loops accumulating k = 0..n-1, first term
//                  initialising (matrix.h:43-60) — so host-side plant updates round like the reference's.
// It is deliberately NOT an expression-template library: everything is evaluated eagerly into a value.
#pragma once
#include <cstddef>

namespace bmpc {

template <typename T, long NR_, long NC_>
class matrix {
    static_assert(NR_ > 0 && NC_ > 0, "bmpc::matrix is fixed-size only");
    T d_[NR_ * NC_];

public:
    typedef T type;
    static const long NR = NR_;
    static const long NC = NC_;

    // dlib leaves fixed-size storage uninitialised (matrix_data_layout.h:137); zero-filling is a valid refinement
    matrix() { for (long k = 0; k < NR_ * NC_; ++k) d_[k] = T(); }

    long nr() const { return NR_; }
    long nc() const { return NC_; }
    long size() const { return NR_ * NC_; }
    // dlib's dynamically sized matrices have set_size(); on a fixed-size one it must match (mpc.h:42-45 calls it)
    void set_size(long, long = 1) {}

    T& operator()(long r, long c) { return d_[r * NC_ + c]; }
    const T& operator()(long r, long c) const { return d_[r * NC_ + c]; }
    T& operator()(long i) { return d_[i]; }                  // vectors (and flat row-major access)
    const T& operator()(long i) const { return d_[i]; }
    T* data() { return d_; }
    const T* data() const { return d_; }

    // `m = 1, 2, 3, 4;`
    class comma_filler {
        matrix* m_;
        long next_;
    public:
        comma_filler(matrix* m, long next) : m_(m), next_(next) {}
        comma_filler operator,(const T& v) const
        {
            if (next_ < NR_ * NC_) m_->d_[next_] = v;
            return comma_filler(m_, next_ + 1);
        }
    };
    comma_filler operator=(const T& v)
    {
        for (long k = 0; k < NR_ * NC_; ++k) d_[k] = v;     // a lone scalar fills the matrix
        return comma_filler(this, 1);
    }

    matrix& operator+=(const matrix& o) { for (long k = 0; k < NR_ * NC_; ++k) d_[k] = d_[k] + o.d_[k]; return *this; }
    matrix& operator-=(const matrix& o) { for (long k = 0; k < NR_ * NC_; ++k) d_[k] = d_[k] - o.d_[k]; return *this; }
};

template <typename T, long NR, long NC>
matrix<T, NR, NC> operator+(const matrix<T, NR, NC>& a, const matrix<T, NR, NC>& b)
{
    matrix<T, NR, NC> r;
    for (long k = 0; k < NR * NC; ++k) r(k) = a(k) + b(k);
    return r;
}
template <typename T, long NR, long NC>
matrix<T, NR, NC> operator-(const matrix<T, NR, NC>& a, const matrix<T, NR, NC>& b)
{
    matrix<T, NR, NC> r;
    for (long k = 0; k < NR * NC; ++k) r(k) = a(k) - b(k);
    return r;
}
template <typename T, long NR, long NC>
matrix<T, NR, NC> operator-(const matrix<T, NR, NC>& a)
{
    matrix<T, NR, NC> r;
    for (long k = 0; k < NR * NC; ++k) r(k) = -a(k);
    return r;
}
template <typename T, long NR, long NK, long NC>
matrix<T, NR, NC> operator*(const matrix<T, NR, NK>& a, const matrix<T, NK, NC>& b)
{
    matrix<T, NR, NC> r;
    for (long i = 0; i < NR; ++i)
        for (long j = 0; j < NC; ++j) {
            T t = a(i, 0) * b(0, j);
            for (long k = 1; k < NK; ++k) t += a(i, k) * b(k, j);
            r(i, j) = t;
        }
    return r;
}
template <typename T, long NR, long NC>
matrix<T, NR, NC> operator*(const matrix<T, NR, NC>& a, const T& s)
{
    matrix<T, NR, NC> r;
    for (long k = 0; k < NR * NC; ++k) r(k) = a(k) * s;
    return r;
}
template <typename T, long NR, long NC>
matrix<T, NR, NC> operator*(const T& s, const matrix<T, NR, NC>& a)
{
    matrix<T, NR, NC> r;
    for (long k = 0; k < NR * NC; ++k) r(k) = s * a(k);
    return r;
}
template <typename T, long NR, long NC>
matrix<T, NC, NR> trans(const matrix<T, NR, NC>& a)
{
    matrix<T, NC, NR> r;
    for (long i = 0; i < NR; ++i)
        for (long j = 0; j < NC; ++j) r(j, i) = a(i, j);
    return r;
}

}  // namespace bmpc
