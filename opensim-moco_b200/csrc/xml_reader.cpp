#include "xml_reader.h"

#include <cctype>
#include <cstdlib>
#include <fstream>
#include <sstream>

namespace mocob200 {

std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && std::isspace((unsigned char)s[a])) ++a;
    while (b > a && std::isspace((unsigned char)s[b - 1])) --b;
    return s.substr(a, b - a);
}

std::vector<double> parseDoubles(const std::string& s) {
    std::vector<double> v;
    const char* p = s.c_str();
    char* e = nullptr;
    while (*p) {
        while (*p && (std::isspace((unsigned char)*p) || *p == ',')) ++p;
        if (!*p) break;
        double d = std::strtod(p, &e);
        if (e == p) throw std::runtime_error("parseDoubles: bad number in '" + s + "'");
        v.push_back(d);
        p = e;
    }
    return v;
}

std::string XmlNode::childText(const std::string& t, const std::string& def) const {
    const XmlNode* c = child(t);
    return c ? trim(c->text) : def;
}

namespace {
struct Parser {
    const std::string& s;
    size_t i = 0;
    explicit Parser(const std::string& str) : s(str) {}
    [[noreturn]] void fail(const std::string& m) {
        throw std::runtime_error("xml parse error at byte " + std::to_string(i) + ": " + m);
    }
    bool starts(const char* lit) const { return s.compare(i, std::char_traits<char>::length(lit), lit) == 0; }
    void skipWs() {
        while (i < s.size() && std::isspace((unsigned char)s[i])) ++i;
    }
    void skipMisc() { // whitespace, comments, declarations, doctype
        for (;;) {
            skipWs();
            if (starts("<!--")) {
                size_t e = s.find("-->", i);
                if (e == std::string::npos) fail("unterminated comment");
                i = e + 3;
            } else if (starts("<?")) {
                size_t e = s.find("?>", i);
                if (e == std::string::npos) fail("unterminated declaration");
                i = e + 2;
            } else if (starts("<!")) {
                size_t e = s.find('>', i);
                if (e == std::string::npos) fail("unterminated <!");
                i = e + 1;
            } else
                return;
        }
    }
    static std::string unescape(const std::string& t) {
        if (t.find('&') == std::string::npos) return t;
        std::string o;
        for (size_t k = 0; k < t.size(); ++k) {
            if (t[k] == '&') {
                if (!t.compare(k, 4, "&lt;")) { o += '<'; k += 3; }
                else if (!t.compare(k, 4, "&gt;")) { o += '>'; k += 3; }
                else if (!t.compare(k, 5, "&amp;")) { o += '&'; k += 4; }
                else if (!t.compare(k, 6, "&quot;")) { o += '"'; k += 5; }
                else if (!t.compare(k, 6, "&apos;")) { o += '\''; k += 5; }
                else o += t[k];
            } else
                o += t[k];
        }
        return o;
    }
    std::string parseName() {
        size_t a = i;
        while (i < s.size() && (std::isalnum((unsigned char)s[i]) || s[i] == '_' || s[i] == '-' || s[i] == ':' || s[i] == '.')) ++i;
        if (i == a) fail("expected a name");
        return s.substr(a, i - a);
    }
    std::unique_ptr<XmlNode> parseElement() {
        if (s[i] != '<') fail("expected '<'");
        ++i;
        auto n = std::make_unique<XmlNode>();
        n->tag = parseName();
        for (;;) {
            skipWs();
            if (i >= s.size()) fail("eof in tag");
            if (s[i] == '/') {
                if (s[i + 1] != '>') fail("bad empty-element tag");
                i += 2;
                return n;
            }
            if (s[i] == '>') { ++i; break; }
            std::string k = parseName();
            skipWs();
            if (s[i] != '=') fail("expected '='");
            ++i;
            skipWs();
            char q = s[i];
            if (q != '"' && q != '\'') fail("expected quote");
            size_t e = s.find(q, i + 1);
            if (e == std::string::npos) fail("unterminated attribute");
            n->attrs.emplace_back(k, unescape(s.substr(i + 1, e - i - 1)));
            i = e + 1;
        }
        // content
        for (;;) {
            size_t lt = s.find('<', i);
            if (lt == std::string::npos) fail("eof in element " + n->tag);
            n->text += unescape(s.substr(i, lt - i));
            i = lt;
            if (starts("<!--")) {
                size_t e = s.find("-->", i);
                if (e == std::string::npos) fail("unterminated comment");
                i = e + 3;
            } else if (starts("<![CDATA[")) {
                size_t e = s.find("]]>", i);
                if (e == std::string::npos) fail("unterminated CDATA");
                n->text += s.substr(i + 9, e - i - 9);
                i = e + 3;
            } else if (starts("</")) {
                i += 2;
                std::string t = parseName();
                if (t != n->tag) fail("mismatched close tag " + t + " for " + n->tag);
                skipWs();
                if (s[i] != '>') fail("expected '>'");
                ++i;
                return n;
            } else {
                n->children.push_back(parseElement());
            }
        }
    }
};
} // namespace

std::unique_ptr<XmlNode> parseXmlString(const std::string& str) {
    Parser p(str);
    p.skipMisc();
    if (p.i >= str.size()) throw std::runtime_error("xml: empty document");
    return p.parseElement();
}

std::unique_ptr<XmlNode> parseXmlFile(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open '" + path + "'");
    std::stringstream ss;
    ss << f.rdbuf();
    return parseXmlString(ss.str());
}

} // namespace mocob200
