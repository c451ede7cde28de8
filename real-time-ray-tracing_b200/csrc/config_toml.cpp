// rtrt_load_config_toml / rtrt_config_defaults — the reference's LoadConfig (src/configLoader.cpp:5-27)
// with the path as a parameter.  Same tables and keys ([resolution], [file], [optimziation] — sic),
// same defaults (toml::find_or fallbacks).  The reference parses with toml11; this is a small
// hand-rolled parser for the subset config.toml uses: tables, integers, floats, booleans, basic
// strings and arrays of strings, '#' comments.
#include "../../include/rtrt.h"
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>
#include <cctype>

namespace {

std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && isspace((unsigned char)s[a])) ++a;
    while (b > a && isspace((unsigned char)s[b - 1])) --b;
    return s.substr(a, b - a);
}

// strip a trailing comment that is not inside a string
std::string strip_comment(const std::string& s) {
    bool in = false;
    for (size_t i = 0; i < s.size(); ++i) {
        if (s[i] == '"' && (i == 0 || s[i - 1] != '\\')) in = !in;
        if (s[i] == '#' && !in) return s.substr(0, i);
    }
    return s;
}

bool parse_string(const std::string& v, std::string& out) {
    if (v.size() < 2 || v.front() != '"' || v.back() != '"') return false;
    out.clear();
    for (size_t i = 1; i + 1 < v.size(); ++i) {
        if (v[i] == '\\' && i + 2 < v.size()) {
            char c = v[++i];
            out += (c == 'n') ? '\n' : (c == 't') ? '\t' : c;
        } else out += v[i];
    }
    return true;
}

bool parse_string_array(const std::string& v, std::vector<std::string>& out) {
    if (v.size() < 2 || v.front() != '[' || v.back() != ']') return false;
    out.clear();
    std::string body = v.substr(1, v.size() - 2), cur;
    bool in = false;
    for (size_t i = 0; i <= body.size(); ++i) {
        char ch = i < body.size() ? body[i] : ',';
        if (ch == '"' && (i == 0 || body[i - 1] != '\\')) in = !in;
        if (ch == ',' && !in) {
            std::string t = trim(cur), s;
            if (!t.empty()) { if (!parse_string(t, s)) return false; out.push_back(s); }
            cur.clear();
        } else cur += ch;
    }
    return true;
}

void copy_str(char* dst, size_t cap, const std::string& s) {
    std::snprintf(dst, cap, "%s", s.c_str());
}

}  // namespace

extern "C" {

int rtrt_config_defaults(rtrt_config* c) {
    if (!c) return RTRT_ERR_ARG;
    std::memset(c, 0, sizeof(*c));
    c->width = 1920; c->height = 1080;                  // src/configLoader.cpp:13-14
    c->loadCameraAtInit = 0;                             // :20
    c->useDynamicResolution = 1; c->targetFps = 60.0f;   // :22-23
    c->maxWidth = 3840; c->maxHeight = 2160;             // :24-25
    c->minWidth = 640; c->minHeight = 480;               // :26-27
    return RTRT_OK;
}

int rtrt_load_config_toml(const char* path, rtrt_config* c) {
    if (!path || !c) return RTRT_ERR_ARG;
    rtrt_config_defaults(c);
    FILE* f = std::fopen(path, "r");
    if (!f) return RTRT_ERR_IO;
    std::string table;
    char line[4096];
    int rc = RTRT_OK;
    while (std::fgets(line, sizeof(line), f)) {
        std::string s = trim(strip_comment(line));
        if (s.empty()) continue;
        if (s.front() == '[' && s.back() == ']' && s.find('=') == std::string::npos) { table = trim(s.substr(1, s.size() - 2)); continue; }
        size_t eq = s.find('=');
        if (eq == std::string::npos) { rc = RTRT_ERR_IO; break; }
        const std::string key = trim(s.substr(0, eq)), val = trim(s.substr(eq + 1));
        std::string sv; std::vector<std::string> av;
        auto as_int = [&](int32_t& dst) { char* e; long v = std::strtol(val.c_str(), &e, 10); if (e == val.c_str()) rc = RTRT_ERR_IO; else dst = (int32_t)v; };
        auto as_bool = [&](int32_t& dst) { if (val == "true") dst = 1; else if (val == "false") dst = 0; else rc = RTRT_ERR_IO; };
        if (table == "resolution") {
            if (key == "width") as_int(c->width);
            else if (key == "height") as_int(c->height);
        } else if (table == "file") {
            if (key == "inputMeshFileName") { if (parse_string(val, sv)) copy_str(c->inputMeshFileName, sizeof(c->inputMeshFileName), sv); else rc = RTRT_ERR_IO; }
            else if (key == "inputCameraFileName") { if (parse_string(val, sv)) copy_str(c->inputCameraFileName, sizeof(c->inputCameraFileName), sv); else rc = RTRT_ERR_IO; }
            else if (key == "cameraSaveFileName") { if (parse_string(val, sv)) copy_str(c->cameraSaveFileName, sizeof(c->cameraSaveFileName), sv); else rc = RTRT_ERR_IO; }
            else if (key == "loadCameraAtInit") as_bool(c->loadCameraAtInit);
            else if (key == "inputTextureFileNames") {
                if (!parse_string_array(val, av)) rc = RTRT_ERR_IO;
                else {
                    c->numInputTextureFileNames = (int32_t)(av.size() < 4 ? av.size() : 4);
                    for (int i = 0; i < c->numInputTextureFileNames; ++i) copy_str(c->inputTextureFileNames[i], 256, av[i]);
                }
            }
        } else if (table == "optimziation") {           // sic — src/configLoader.cpp:11
            if (key == "useDynamicResolution") as_bool(c->useDynamicResolution);
            else if (key == "targetFps") { char* e; double v = std::strtod(val.c_str(), &e); if (e == val.c_str()) rc = RTRT_ERR_IO; else c->targetFps = (float)v; }
            else if (key == "maxWidth") as_int(c->maxWidth);
            else if (key == "maxHeight") as_int(c->maxHeight);
            else if (key == "minWidth") as_int(c->minWidth);
            else if (key == "minHeight") as_int(c->minHeight);
        }
        if (rc) break;
    }
    std::fclose(f);
    return rc;
}

}  // extern "C"
