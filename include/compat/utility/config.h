// aniso_b200 drop-in header: `key = value` configuration files (utility/config.{h,cpp}).  Lines need spaces around '=';
// '#' starts a comment.  Unlike the reference on libstdc++ >= 13 a value at end of line is accepted
// with or without a trailing space.
#ifndef ANISO_B200_COMPAT_CONFIG_H
#define ANISO_B200_COMPAT_CONFIG_H

#include <map>
#include <sstream>

#include "../bbfmm/utils.h"

class config {
public:
    config() {}
    ~config() {}
    config(char* path) {
        std::ifstream f(path, std::ifstream::in);
        parse(f);
    }
    std::map<std::string, std::string> options;
    void parse(std::istream& cfgFile) {
        for (std::string line; std::getline(cfgFile, line);) {
            std::istringstream iss(line);
            std::string id, eq, val, extra;
            if (!(iss >> id)) continue;      // blank line
            if (id[0] == '#') continue;
            if (!(iss >> eq >> val) || eq != "=" || (iss >> extra)) {
                std::cout << "Incomplete configuration! Did you forget the space?" << std::endl;
                continue;
            }
            options[id] = val;
        }
    }
    void print() {
        std::cout << std::setw(15) << "=============== OPTIONS ===============" << std::endl;
        for (auto& kv : options) std::cout << std::setw(15) << kv.first << " = " << kv.second << std::endl;
        std::cout << std::setw(15) << "=======================================" << std::endl;
    }
};

#endif
