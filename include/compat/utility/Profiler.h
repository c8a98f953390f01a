// aniso_b200 drop-in header: wall-clock tic/toc table with the interface of utility/Profiler.h:13-68.
#ifndef ANISO_B200_COMPAT_PROFILER_H
#define ANISO_B200_COMPAT_PROFILER_H

#include "../bbfmm/utils.h"

class Profiler {
public:
    Profiler() {}
    ~Profiler() {
        double total = 0;
        for (auto& kv : acc_) total += kv.second;
        std::cout << std::setw(15) << "=============== Profiler ===============" << std::endl;
        for (auto& name : order_)
            std::cout << std::fixed << std::setprecision(3) << std::setw(30) << name << "[C]" << std::setw(15)
                      << (total > 0 ? acc_[name] / total * 100 : 0.0) << "%" << std::setw(15) << acc_[name] << " seconds"
                      << std::endl;
        std::cout << std::setw(30) << "counted time" << std::setw(34) << total << " seconds" << std::endl;
    }
    void tic(std::string s = "") {
        cur_ = s;
        begin_ = std::chrono::steady_clock::now();
    }
    void toc(bool count = true) {
        const double dt = std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - begin_).count() / 1e6;
        if (!count) return;
        if (!acc_.count(cur_)) order_.push_back(cur_);
        acc_[cur_] += dt;
    }

private:
    std::chrono::steady_clock::time_point begin_;
    std::string cur_;
    unordered_map<std::string, double> acc_;
    vector<std::string> order_;
};

#endif
