// aniso_b200 drop-in header: CSV writers / reader with the formats of matlab_io.h:14-56 (32-digit scientific).
#ifndef ANISO_B200_COMPAT_MATLAB_IO_H
#define ANISO_B200_COMPAT_MATLAB_IO_H

#include "bbfmm/bbfmm.h"
#include "bbfmm/linalg.h"
#include "bbfmm/utils.h"

inline void write_to_csv(vector<scalar_t>& data, std::string filename) {
    std::ofstream out(filename);
    for (auto v : data) out << v << "\n";
}
inline void write_to_csv(bbfmm::Vector& data, std::string filename) {
    std::ofstream out(filename);
    for (int i = 0; i < data.row(); ++i) out << std::scientific << std::setprecision(32) << std::setw(32) << data(i) << "\n";
}
inline void write_to_csv(vector<bbfmm::point>& data, std::string filename, std::string sep = " ") {
    std::ofstream out(filename);
    for (auto& p : data)
        out << std::scientific << std::setprecision(32) << std::setw(32) << p.x << sep << std::setw(32) << p.y << "\n";
}
// warm start: reads as many values as the file holds; a missing file leaves `data` untouched (main.cpp:140)
inline void load_csv(bbfmm::Vector& data, std::string filename) {
    std::ifstream in(filename);
    if (!in.is_open()) std::cout << "File not open." << std::endl;
    int row = 0;
    double num = 0.;
    while (in >> num && row < data.row()) data(row++) = num;
}

#endif
