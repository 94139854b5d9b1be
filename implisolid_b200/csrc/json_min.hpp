// Minimal JSON document for the two strings the geometry API receives.
//
// Replaces Boost.PropertyTree's read_json as used by the reference
// (object_factory.hpp:741-758, polygoniser_settings.hpp:147-163).  PropertyTree semantics kept:
// scalars are stored as text and converted on request; a conversion that does not consume the
// whole text fails and the caller's default is used; arrays are children without a key;
// "a.b" paths descend through objects.  Malformed input is fatal (the reference dies on an
// uncaught exception there).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace isd {

class JsonDoc {
public:
    struct Item {
        uint32_t key_off, key_len;     // into text_ (0,0 for array elements / root)
        uint32_t val_off, val_len;     // scalar text (containers: 0,0)
        int32_t first_child, next_sibling;
        char kind;                     // 'o' object, 'a' array, 's' scalar
    };

    explicit JsonDoc(const char* src) {
        p_ = src;
        items_.reserve(64);
        int root = parse_value(0, 0);
        skip_ws();
        if (*p_) die("trailing characters");
        (void)root;
    }

    int root() const { return 0; }
    const Item& item(int i) const { return items_[i]; }

    int child(int parent, const char* key) const {
        size_t kl = strlen(key);
        for (int c = items_[parent].first_child; c >= 0; c = items_[c].next_sibling)
            if (items_[c].key_len == kl && memcmp(text_.data() + items_[c].key_off, key, kl) == 0) return c;
        return -1;
    }
    // dotted path lookup, -1 when absent
    int find(int from, const char* dotted) const {
        int cur = from;
        const char* s = dotted;
        while (cur >= 0) {
            const char* dot = strchr(s, '.');
            std::string key = dot ? std::string(s, dot) : std::string(s);
            cur = child(cur, key.c_str());
            if (!dot) break;
            s = dot + 1;
        }
        return cur;
    }
    int first_child(int i) const { return items_[i].first_child; }
    int next(int i) const { return items_[i].next_sibling; }
    int count_children(int i) const {
        int n = 0;
        for (int c = items_[i].first_child; c >= 0; c = items_[c].next_sibling) ++n;
        return n;
    }
    std::string text(int i) const { return std::string(text_.data() + items_[i].val_off, items_[i].val_len); }

    bool to_float(int i, float* out) const {
        if (i < 0 || items_[i].kind != 's' || items_[i].val_len == 0) return false;
        std::string t = text(i);
        char* end = nullptr;
        float v = strtof(t.c_str(), &end);
        if (end == t.c_str() || !only_ws(end)) return false;
        *out = v;
        return true;
    }
    bool to_int(int i, int* out) const {
        if (i < 0 || items_[i].kind != 's' || items_[i].val_len == 0) return false;
        std::string t = text(i);
        char* end = nullptr;
        long v = strtol(t.c_str(), &end, 10);
        if (end == t.c_str() || !only_ws(end)) return false;
        *out = (int)v;
        return true;
    }
    bool to_bool(int i, bool* out) const {
        if (i < 0 || items_[i].kind != 's') return false;
        std::string t = text(i);
        if (t == "1" || t == "true") { *out = true; return true; }
        if (t == "0" || t == "false") { *out = false; return true; }
        return false;
    }
    float get_float(const char* path, float dflt) const { float v; return to_float(find(0, path), &v) ? v : dflt; }
    int get_int(const char* path, int dflt) const { int v; return to_int(find(0, path), &v) ? v : dflt; }
    bool get_bool(const char* path, bool dflt) const { bool v; return to_bool(find(0, path), &v) ? v : dflt; }

private:
    const char* p_;
    std::vector<Item> items_;
    std::vector<char> text_;

    static bool only_ws(const char* s) {
        while (*s == ' ' || *s == '\t' || *s == '\n' || *s == '\r') ++s;
        return *s == 0;
    }
    [[noreturn]] void die(const char* why) const {
        fprintf(stderr, "implisolid_b200: JSON parse error: %s near \"%.24s\"\n", why, p_);
        abort();
    }
    void skip_ws() { while (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r') ++p_; }

    uint32_t stash(const std::string& s) {
        uint32_t off = (uint32_t)text_.size();
        text_.insert(text_.end(), s.begin(), s.end());
        return off;
    }
    std::string parse_string() {
        std::string out;
        ++p_;
        for (;;) {
            char c = *p_;
            if (!c) die("unterminated string");
            if (c == '"') { ++p_; break; }
            if (c == '\\') {
                c = *++p_;
                switch (c) {
                    case 'n': out.push_back('\n'); break;
                    case 't': out.push_back('\t'); break;
                    case 'r': out.push_back('\r'); break;
                    case 'b': out.push_back('\b'); break;
                    case 'f': out.push_back('\f'); break;
                    case 'u': {
                        unsigned cp = 0;
                        for (int k = 0; k < 4; k++) {
                            char h = *++p_;
                            cp <<= 4;
                            if (h >= '0' && h <= '9') cp |= (unsigned)(h - '0');
                            else if (h >= 'a' && h <= 'f') cp |= (unsigned)(h - 'a' + 10);
                            else if (h >= 'A' && h <= 'F') cp |= (unsigned)(h - 'A' + 10);
                            else die("bad \\u escape");
                        }
                        if (cp < 0x80) out.push_back((char)cp);
                        else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 63))); }
                        else { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 63))); out.push_back((char)(0x80 | (cp & 63))); }
                        break;
                    }
                    case 0: die("unterminated escape");
                    default: out.push_back(c);
                }
                ++p_;
            } else {
                out.push_back(c);
                ++p_;
            }
        }
        return out;
    }
    int new_item(uint32_t koff, uint32_t klen, char kind) {
        Item it;
        it.key_off = koff; it.key_len = klen; it.val_off = 0; it.val_len = 0;
        it.first_child = -1; it.next_sibling = -1; it.kind = kind;
        items_.push_back(it);
        return (int)items_.size() - 1;
    }
    int parse_value(uint32_t koff, uint32_t klen) {
        skip_ws();
        if (*p_ == '{' || *p_ == '[') {
            const bool obj = (*p_ == '{');
            const char close = obj ? '}' : ']';
            int self = new_item(koff, klen, obj ? 'o' : 'a');
            ++p_;
            skip_ws();
            if (*p_ == close) { ++p_; return self; }
            int last = -1;
            for (;;) {
                uint32_t ko = 0, kl = 0;
                if (obj) {
                    skip_ws();
                    if (*p_ != '"') die("expected member name");
                    std::string k = parse_string();
                    ko = stash(k); kl = (uint32_t)k.size();
                    skip_ws();
                    if (*p_ != ':') die("expected ':'");
                    ++p_;
                }
                int c = parse_value(ko, kl);
                if (last < 0) items_[self].first_child = c; else items_[last].next_sibling = c;
                last = c;
                skip_ws();
                if (*p_ == ',') { ++p_; continue; }
                if (*p_ == close) { ++p_; break; }
                die("expected ',' or closing bracket");
            }
            return self;
        }
        int self = new_item(koff, klen, 's');
        std::string v;
        if (*p_ == '"') {
            v = parse_string();
        } else {
            const char* b = p_;
            while (*p_ && !strchr(",}] \t\r\n", *p_)) ++p_;
            if (p_ == b) die("expected a value");
            v.assign(b, p_);
        }
        items_[self].val_off = stash(v);
        items_[self].val_len = (uint32_t)v.size();
        return self;
    }
};

}  // namespace isd
