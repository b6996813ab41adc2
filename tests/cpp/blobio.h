// Tiny named-array container shared with tests/cppio.py: per array one text line
// "<name> <dtype> <count>\n" followed by the raw little-endian bytes.
#pragma once
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

struct Blob {
    std::map<std::string, std::vector<char> > a;
    std::map<std::string, std::string> dt;
    static size_t esz(const std::string& d) { return d == "u1" ? 1 : 4; }
    void load(const char* path) {
        FILE* f = fopen(path, "rb");
        if (!f) throw std::runtime_error(std::string("cannot open ") + path);
        char name[128], d[16]; long long n;
        while (fscanf(f, "%127s %15s %lld", name, d, &n) == 3) {
            fgetc(f);  // newline
            std::vector<char> v((size_t)n * esz(d));
            if (!v.empty() && fread(v.data(), 1, v.size(), f) != v.size()) throw std::runtime_error("short read");
            a[name] = v; dt[name] = d;
        }
        fclose(f);
    }
    template <class T> const T* get(const std::string& k) const {
        std::map<std::string, std::vector<char> >::const_iterator it = a.find(k);
        if (it == a.end()) throw std::runtime_error("missing array " + k);
        return reinterpret_cast<const T*>(it->second.data());
    }
    size_t count(const std::string& k) const { return a.at(k).size() / esz(dt.at(k)); }
    bool has(const std::string& k) const { return a.count(k) != 0; }
};

struct BlobWriter {
    FILE* f;
    explicit BlobWriter(const char* path) : f(fopen(path, "wb")) { if (!f) throw std::runtime_error("cannot write"); }
    ~BlobWriter() { if (f) fclose(f); }
    void put(const char* name, const char* dtype, const void* p, size_t count) {
        fprintf(f, "%s %s %lld\n", name, dtype, (long long)count);
        if (count) fwrite(p, Blob::esz(dtype), count, f);
    }
};
