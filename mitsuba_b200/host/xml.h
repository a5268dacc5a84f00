// xml.h -- minimal XML reader for Mitsuba scene files (replaces the xerces-c SAX parser used by
// the reference, src/librender/scenehandler.cpp:197-820).  Supports what scene files use:
// prolog, comments, elements with quoted attributes, self-closing tags, character entities.
// Text content is ignored (scene files carry all data in attributes).
#pragma once
#include <cctype>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace mtsbhost {

struct XmlNode {
    std::string tag;
    std::map<std::string, std::string> attr;
    std::vector<std::unique_ptr<XmlNode>> children;
    int line = 0;
    bool has(const std::string &k) const { return attr.find(k) != attr.end(); }
    std::string get(const std::string &k, const std::string &dflt = "") const {
        auto it = attr.find(k);
        return it == attr.end() ? dflt : it->second;
    }
};

class XmlParser {
public:
    XmlParser(const std::string &text, const std::string &name) : s(text), file(name) {}

    std::unique_ptr<XmlNode> parse() {
        skipMisc();
        std::unique_ptr<XmlNode> root = element();
        skipMisc();
        if (pos < s.size()) fail("content after the root element");
        return root;
    }

private:
    const std::string &s;
    std::string file;
    size_t pos = 0;
    int line = 1;

    [[noreturn]] void fail(const std::string &msg) const {
        throw std::runtime_error("Error while parsing \"" + file + "\" (near line " + std::to_string(line) + "): " + msg);
    }
    bool startsWith(const char *t) const { return s.compare(pos, strlen(t), t) == 0; }
    void advance(size_t n) { for (size_t i = 0; i < n && pos < s.size(); ++i) { if (s[pos] == '\n') ++line; ++pos; } }
    void skipWs() { while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\t' || s[pos] == '\r' || s[pos] == '\n')) advance(1); }
    void skipUntil(const char *t) {
        size_t e = s.find(t, pos);
        if (e == std::string::npos) fail(std::string("unterminated construct, expected ") + t);
        advance(e - pos + strlen(t));
    }
    void skipMisc() {
        for (;;) {
            skipWs();
            if (startsWith("<?")) skipUntil("?>");
            else if (startsWith("<!--")) skipUntil("-->");
            else if (startsWith("<!")) skipUntil(">");
            else break;
        }
    }
    static bool nameChar(char c) { return isalnum((unsigned char) c) || c == '_' || c == '-' || c == ':' || c == '.'; }
    std::string name() {
        size_t b = pos;
        while (pos < s.size() && nameChar(s[pos])) ++pos;
        if (b == pos) fail("expected a name");
        return s.substr(b, pos - b);
    }
    std::string decode(const std::string &v) const {
        std::string o;
        for (size_t i = 0; i < v.size(); ++i) {
            if (v[i] != '&') { o += v[i]; continue; }
            size_t e = v.find(';', i);
            if (e == std::string::npos) fail("bad character entity");
            std::string ent = v.substr(i + 1, e - i - 1);
            if (ent == "lt") o += '<'; else if (ent == "gt") o += '>'; else if (ent == "amp") o += '&';
            else if (ent == "quot") o += '"'; else if (ent == "apos") o += '\'';
            else if (!ent.empty() && ent[0] == '#') o += (char) (ent[1] == 'x' ? strtol(ent.c_str() + 2, nullptr, 16) : strtol(ent.c_str() + 1, nullptr, 10));
            else fail("unknown entity &" + ent + ";");
            i = e;
        }
        return o;
    }
    std::unique_ptr<XmlNode> element() {
        if (pos >= s.size() || s[pos] != '<') fail("expected '<'");
        advance(1);
        std::unique_ptr<XmlNode> n(new XmlNode());
        n->line = line;
        n->tag = name();
        for (;;) {
            skipWs();
            if (pos >= s.size()) fail("unterminated tag <" + n->tag + ">");
            if (s[pos] == '/') { advance(1); if (pos >= s.size() || s[pos] != '>') fail("expected '>'"); advance(1); return n; }
            if (s[pos] == '>') { advance(1); break; }
            std::string k = name();
            skipWs();
            if (pos >= s.size() || s[pos] != '=') fail("expected '=' after attribute " + k);
            advance(1); skipWs();
            if (pos >= s.size() || (s[pos] != '"' && s[pos] != '\'')) fail("attribute value must be quoted");
            char q = s[pos]; advance(1);
            size_t e = s.find(q, pos);
            if (e == std::string::npos) fail("unterminated attribute value");
            if (n->attr.count(k)) fail("duplicate attribute " + k);
            n->attr[k] = decode(s.substr(pos, e - pos));
            advance(e - pos + 1);
        }
        for (;;) {
            // text content is skipped
            size_t lt = s.find('<', pos);
            if (lt == std::string::npos) fail("missing </" + n->tag + ">");
            advance(lt - pos);
            if (startsWith("<!--")) { skipUntil("-->"); continue; }
            if (startsWith("<![CDATA[")) { skipUntil("]]>"); continue; }
            if (startsWith("</")) {
                advance(2);
                std::string c = name();
                if (c != n->tag) fail("mismatched closing tag </" + c + "> for <" + n->tag + ">");
                skipWs();
                if (pos >= s.size() || s[pos] != '>') fail("expected '>'");
                advance(1);
                return n;
            }
            n->children.push_back(element());
        }
    }
};

}  // namespace mtsbhost
