// dqmcb/utility.h -- host mirror of the two `utility` types the sweep path's interface mentions
// (reference: include/utility.h:19-48 `utility::random`, :50-276 `utility::parameters`).
//
// `random` must reproduce the reference's draws exactly, so it wraps the same std::mt19937 and the
// same libstdc++ distributions.  `parameters` keeps the reference's typed getters
// (getInt/getDouble/getBool/getString/getDoubleVector, optional defaults, hasSection/hasKey, INI
// sections, `#`/`;` comments, `_` digit separators) and adds programmatic `set()` for embedding.
#pragma once
#include <algorithm>
#include <cctype>
#include <fstream>
#include <map>
#include <random>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace utility {

class random {
   private:
    std::mt19937 generator_;
    std::uniform_int_distribution<int> dist_GHQField_;

   public:
    random(unsigned int seed) : generator_(seed), dist_GHQField_(0, 3) {}
    // include/utility.h:34-37
    bool bernoulli(double p) {
        std::bernoulli_distribution dist(p);
        return dist(generator_);
    }
    int rand_GHQField() { return dist_GHQField_(generator_); }
    std::mt19937& get_generator() { return generator_; }
};

class parameters {
   private:
    std::map<std::string, std::map<std::string, std::string>> data_;

    static std::string trim(const std::string& s) {
        size_t a = 0, b = s.size();
        while (a < b && std::isspace((unsigned char)s[a])) ++a;
        while (b > a && std::isspace((unsigned char)s[b - 1])) --b;
        return s.substr(a, b - a);
    }
    static std::string strip_digit_separators(const std::string& v) {
        std::string o;
        for (char c : v)
            if (c != '_') o.push_back(c);
        return o;
    }

   public:
    parameters() = default;
    explicit parameters(const std::string& filename) { load(filename); }

    void load(const std::string& filename) {
        std::ifstream in(filename);
        if (!in) throw std::runtime_error("Could not open parameter file: " + filename);
        std::string line, section;
        while (std::getline(in, line)) {
            const size_t cpos = line.find_first_of("#;");
            if (cpos != std::string::npos) line = line.substr(0, cpos);
            line = trim(line);
            if (line.empty()) continue;
            if (line.front() == '[' && line.back() == ']') {
                section = trim(line.substr(1, line.size() - 2));
                continue;
            }
            const size_t eq = line.find('=');
            if (eq == std::string::npos) continue;
            data_[section][trim(line.substr(0, eq))] = trim(line.substr(eq + 1));
        }
    }
    // programmatic construction (not in the reference; embedding / tests)
    parameters& set(const std::string& section, const std::string& key, const std::string& value) {
        data_[section][key] = value;
        return *this;
    }
    parameters& set(const std::string& section, const std::string& key, double value) {
        std::ostringstream os;
        os.precision(17);
        os << value;
        return set(section, key, os.str());
    }
    parameters& set(const std::string& section, const std::string& key, int value) { return set(section, key, std::to_string(value)); }

    bool hasSection(const std::string& section) const { return data_.count(section) != 0; }
    bool hasKey(const std::string& section, const std::string& key) const {
        auto it = data_.find(section);
        return it != data_.end() && it->second.count(key) != 0;
    }
    std::string getString(const std::string& section, const std::string& key) const {
        auto it = data_.find(section);
        if (it == data_.end()) throw std::runtime_error("Section not found: " + section);
        auto jt = it->second.find(key);
        if (jt == it->second.end()) throw std::runtime_error("Key not found: " + key + " in section " + section);
        return jt->second;
    }
    std::string getString(const std::string& section, const std::string& key, const std::string& dflt) const {
        return hasKey(section, key) ? getString(section, key) : dflt;
    }
    int getInt(const std::string& section, const std::string& key) const {
        const std::string v = strip_digit_separators(getString(section, key));
        try {
            return std::stoi(v);
        } catch (...) {
            throw std::runtime_error("Invalid integer for " + section + "." + key + ": " + v);
        }
    }
    int getInt(const std::string& section, const std::string& key, int dflt) const { return hasKey(section, key) ? getInt(section, key) : dflt; }
    double getDouble(const std::string& section, const std::string& key) const {
        const std::string v = strip_digit_separators(getString(section, key));
        try {
            return std::stod(v);
        } catch (...) {
            throw std::runtime_error("Invalid double for " + section + "." + key + ": " + v);
        }
    }
    double getDouble(const std::string& section, const std::string& key, double dflt) const {
        return hasKey(section, key) ? getDouble(section, key) : dflt;
    }
    bool getBool(const std::string& section, const std::string& key) const {
        std::string v = getString(section, key);
        std::transform(v.begin(), v.end(), v.begin(), [](unsigned char c) { return (char)std::tolower(c); });
        if (v == "true" || v == "1" || v == "yes" || v == "on") return true;
        if (v == "false" || v == "0" || v == "no" || v == "off") return false;
        throw std::runtime_error("Invalid boolean for " + section + "." + key + ": " + v);
    }
    bool getBool(const std::string& section, const std::string& key, bool dflt) const { return hasKey(section, key) ? getBool(section, key) : dflt; }
    std::vector<double> getDoubleVector(const std::string& section, const std::string& key) const {
        std::vector<double> out;
        std::stringstream ss(getString(section, key));
        std::string item;
        while (std::getline(ss, item, ',')) {
            item = trim(item);
            if (!item.empty()) out.push_back(std::stod(strip_digit_separators(item)));
        }
        return out;
    }
};

}  // namespace utility
