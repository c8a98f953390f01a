// aniso_b200 drop-in header: bbfmm::Vector / bbfmm::Matrix with the value semantics of the reference containers
// (bbfmm/linalg.h:11-132, 157-285: deep copies, column-major Matrix, operator()(i) / (i,j), row()/col()/data()).
#ifndef ANISO_B200_COMPAT_LINALG_H
#define ANISO_B200_COMPAT_LINALG_H

#include "utils.h"

namespace bbfmm {

class Vector {
public:
    index_t _row;
    bool_t _ownership;
    scalar_t* _data;

    Vector(int row = 0) : _row(row), _ownership(true), _data(row > 0 ? new scalar_t[row]() : nullptr) {}
    Vector(int row, bool_t ownership, scalar_t* data) : _row(row), _ownership(ownership), _data(nullptr) {
        if (!ownership) {
            _data = data;
        } else if (row > 0) {
            _data = new scalar_t[row];
            std::memcpy(_data, data, sizeof(scalar_t) * row);
        }
    }
    Vector(const Vector& v) : _row(v._row), _ownership(true), _data(v._row > 0 ? new scalar_t[v._row] : nullptr) {
        if (_row > 0) std::memcpy(_data, v._data, sizeof(scalar_t) * _row);
    }
    ~Vector() {
        if (_ownership) delete[] _data;
    }
    Vector& operator=(const Vector& v) {
        if (this == &v) return *this;
        if (_ownership) delete[] _data;
        _row = v._row;
        _ownership = true;
        _data = _row > 0 ? new scalar_t[_row] : nullptr;
        if (_row > 0) std::memcpy(_data, v._data, sizeof(scalar_t) * _row);
        return *this;
    }
    scalar_t* data() { return _data; }
    const scalar_t* data() const { return _data; }
    index_t row() const { return _row; }
    scalar_t& operator()(index_t i) { return _data[i]; }
    const scalar_t& operator()(index_t i) const { return _data[i]; }
    void resize(int row) {
        if (row == _row) return;
        if (_ownership) delete[] _data;
        _row = row;
        _ownership = true;
        _data = row > 0 ? new scalar_t[row]() : nullptr;
    }
};

class Matrix {  // column-major
public:
    index_t _row, _col;
    std::vector<scalar_t> _store;
    Matrix(int row = 0, int col = 0) : _row(row), _col(col), _store((size_t)row * col, 0.0) {}
    scalar_t* data() { return _store.data(); }
    const scalar_t* data() const { return _store.data(); }
    index_t row() const { return _row; }
    index_t col() const { return _col; }
    scalar_t& operator()(index_t i, index_t j) { return _store[(size_t)j * _row + i]; }
    const scalar_t& operator()(index_t i, index_t j) const { return _store[(size_t)j * _row + i]; }
    void resize(int row, int col) {
        _row = row;
        _col = col;
        _store.assign((size_t)row * col, 0.0);
    }
};

inline void setValue(Vector& v, scalar_t val) { std::fill(v.data(), v.data() + v.row(), val); }
inline void setValue(Matrix& m, scalar_t val) { std::fill(m._store.begin(), m._store.end(), val); }

}  // namespace bbfmm

#endif
