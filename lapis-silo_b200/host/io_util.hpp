// File helpers for the dataset builder: whole-file reads, zstd (dlopen of libzstd.so.1, the
// headers are not installed), xz (through the `xz` binary), FASTA and TSV parsing, and a reader
// for the small YAML subset the reference's config files use.
#pragma once
#include <dlfcn.h>

#include <cstdio>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace silo::io {

inline bool fileExists(const std::string& path) { std::ifstream f(path); return f.good(); }

inline std::string readFile(const std::string& path) {
   std::ifstream f(path, std::ios::binary);
   if (!f) throw std::runtime_error("cannot open file " + path);
   std::stringstream ss;
   ss << f.rdbuf();
   return ss.str();
}

inline std::string zstdDecompress(const std::string& data) {
   static void* lib = dlopen("libzstd.so.1", RTLD_NOW);
   if (!lib) throw std::runtime_error("libzstd.so.1 not found");
   using SizeFn = unsigned long long (*)(const void*, size_t);
   using DecompressFn = size_t (*)(void*, size_t, const void*, size_t);
   using IsErrorFn = unsigned (*)(size_t);
   static auto frame_size = reinterpret_cast<SizeFn>(dlsym(lib, "ZSTD_getFrameContentSize"));
   static auto decompress = reinterpret_cast<DecompressFn>(dlsym(lib, "ZSTD_decompress"));
   static auto is_error = reinterpret_cast<IsErrorFn>(dlsym(lib, "ZSTD_isError"));
   unsigned long long n = frame_size(data.data(), data.size());
   if (n == 0ULL - 1 || n == 0ULL - 2) throw std::runtime_error("zstd: unknown frame content size");
   std::string out(n, '\0');
   size_t r = decompress(out.data(), n, data.data(), data.size());
   if (is_error(r)) throw std::runtime_error("zstd: decompression failed");
   out.resize(r);
   return out;
}

inline std::string xzDecompress(const std::string& path) {
   std::string cmd = "xz -dc '" + path + "'";
   FILE* pipe = popen(cmd.c_str(), "r");
   if (!pipe) throw std::runtime_error("cannot run xz");
   std::string out;
   char buf[1 << 16];
   size_t n;
   while ((n = fread(buf, 1, sizeof buf, pipe)) > 0) out.append(buf, n);
   if (pclose(pipe) != 0) throw std::runtime_error("xz failed on " + path);
   return out;
}

// Reads `base`, `base.zst` or `base.xz`, whichever exists; empty optional-like flag via `found`.
inline std::string readMaybeCompressed(const std::string& base, bool& found) {
   found = true;
   if (fileExists(base)) return readFile(base);
   if (fileExists(base + ".zst")) return zstdDecompress(readFile(base + ".zst"));
   if (fileExists(base + ".xz")) return xzDecompress(base + ".xz");
   found = false;
   return "";
}

// common/fasta_reader.cpp:13-46 (FastaReader::nextKey / next): one line with the key after '>', one line with the whole sequence, and
// so on; a key line without '>' (an empty line included) and a key without a following line are FastaFormatExceptions with the
// reference's texts. Lines are what getline yields: split at '\n', nothing trimmed.
struct FastaFormatException : std::runtime_error { using std::runtime_error::runtime_error; };
inline std::vector<std::pair<std::string, std::string>> readFastaEntries(const std::string& text) {
   std::vector<std::pair<std::string, std::string>> out;
   size_t pos = 0;
   auto getline = [&](std::string& line) {      // std::getline on a stream holding `text`
      if (pos >= text.size()) return false;
      size_t eol = text.find('\n', pos);
      if (eol == std::string::npos) eol = text.size();
      line.assign(text, pos, eol - pos);
      pos = eol + 1;
      return true;
   };
   std::string key_with_prefix, genome;
   while (getline(key_with_prefix)) {
      if (key_with_prefix.empty() || key_with_prefix[0] != '>') throw FastaFormatException("Fasta key prefix '>' missing for key: " + key_with_prefix);
      std::string key = key_with_prefix.substr(1);
      if (!getline(genome)) throw FastaFormatException("Missing genome sequence in line following key: " + key);
      out.emplace_back(std::move(key), genome);
   }
   return out;
}
inline std::map<std::string, std::string> parseFasta(const std::string& text) {
   std::map<std::string, std::string> out;
   for (auto& [key, genome] : readFastaEntries(text)) out[key] = std::move(genome);
   return out;
}

struct Table {
   std::vector<std::string> header;
   std::vector<std::vector<std::string>> rows;
   int columnIndex(const std::string& name) const {
      for (size_t k = 0; k < header.size(); ++k) if (header[k] == name) return static_cast<int>(k);
      return -1;
   }
};

inline Table parseTsv(const std::string& text) {
   Table t;
   size_t pos = 0;
   bool first = true;
   while (pos < text.size()) {
      size_t eol = text.find('\n', pos);
      if (eol == std::string::npos) eol = text.size();
      std::string line = text.substr(pos, eol - pos);
      if (!line.empty() && line.back() == '\r') line.pop_back();
      pos = eol + 1;
      if (line.empty()) continue;
      std::vector<std::string> cells;
      size_t s = 0;
      while (true) {
         size_t tab = line.find('\t', s);
         cells.push_back(line.substr(s, tab == std::string::npos ? std::string::npos : tab - s));
         if (tab == std::string::npos) break;
         s = tab + 1;
      }
      if (first) { t.header = cells; first = false; }
      else { cells.resize(t.header.size()); t.rows.push_back(std::move(cells)); }
   }
   return t;
}

// YAML subset: nested maps by indentation, "- " list items of maps, scalar values (optionally quoted).
// Returns flattened entries: for database_config.yaml we only need schema.* scalars and the
// metadata list, so the parser yields (indent, is_list_item, key, value) records.
struct YamlLine { int indent; bool list_item; std::string key; std::string value; };
inline std::vector<YamlLine> parseYamlLines(const std::string& text) {
   std::vector<YamlLine> out;
   std::istringstream in(text);
   std::string line;
   while (std::getline(in, line)) {
      if (!line.empty() && line.back() == '\r') line.pop_back();
      size_t hash = line.find(" #");
      if (hash != std::string::npos) line = line.substr(0, hash);
      size_t k = 0;
      while (k < line.size() && line[k] == ' ') ++k;
      if (k == line.size() || line[k] == '#') continue;
      YamlLine y;
      y.indent = static_cast<int>(k);
      y.list_item = false;
      if (line.compare(k, 2, "- ") == 0) { y.list_item = true; k += 2; y.indent += 2; }
      size_t colon = line.find(':', k);
      if (colon == std::string::npos) continue;
      y.key = line.substr(k, colon - k);
      std::string v = line.substr(colon + 1);
      size_t b = v.find_first_not_of(' ');
      v = b == std::string::npos ? "" : v.substr(b);
      while (!v.empty() && v.back() == ' ') v.pop_back();
      if (v.size() >= 2 && ((v.front() == '"' && v.back() == '"') || (v.front() == '\'' && v.back() == '\''))) v = v.substr(1, v.size() - 2);
      y.value = v;
      out.push_back(y);
   }
   return out;
}

}  // namespace silo::io
