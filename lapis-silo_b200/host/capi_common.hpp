// Shared helpers of the host C entry points (error slot, string duplication, exception mapping).
#pragma once
#include <cstdlib>
#include <cstring>
#include <string>

#include "operators.hpp"

namespace silo::capi {

inline std::string& lastError() { thread_local std::string error; return error; }

inline char* dupString(const std::string& s) {
   char* out = static_cast<char*>(malloc(s.size() + 1));
   memcpy(out, s.c_str(), s.size() + 1);
   return out;
}

// 0 ok; 1 QueryParseException (HTTP 400 in the reference, query_handler.cpp:42-46);
// 2 QueryCompilationException, 3 anything else (HTTP 500, query_handler.cpp:47-73)
template <typename F>
int guarded(F&& f) {
   try { f(); return 0; }
   catch (const QueryParseException& e) { lastError() = std::string("parse:") + e.what(); return 1; }
   catch (const QueryCompilationException& e) { lastError() = std::string("compile:") + e.what(); return 2; }
   catch (const std::exception& e) { lastError() = std::string("other:") + e.what(); return 3; }
}

}  // namespace silo::capi
