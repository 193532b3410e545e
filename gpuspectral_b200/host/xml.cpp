#include "xml.h"

#include <cctype>
#include <stdexcept>

namespace gsbhost {
namespace {

struct Cursor {
    const std::string& s;
    size_t i = 0;
    explicit Cursor(const std::string& text) : s(text) {}
    bool eof() const { return i >= s.size(); }
    bool startsWith(const char* lit) const { return s.compare(i, std::char_traits<char>::length(lit), lit) == 0; }
    void skipSpace() {
        while (!eof() && std::isspace((unsigned char)s[i])) i++;
    }
    [[noreturn]] void fail(const std::string& what) const {
        size_t line = 1;
        for (size_t k = 0; k < i && k < s.size(); k++)
            if (s[k] == '\n') line++;
        throw std::runtime_error("xml: " + what + " at line " + std::to_string(line));
    }
    void skipUntil(const char* lit) {
        size_t p = s.find(lit, i);
        if (p == std::string::npos) fail(std::string("unterminated construct, expected ") + lit);
        i = p + std::char_traits<char>::length(lit);
    }
};

bool isNameChar(char c) { return std::isalnum((unsigned char)c) || c == '_' || c == '-' || c == ':' || c == '.'; }

std::string decodeEntities(const std::string& in) {
    std::string out;
    out.reserve(in.size());
    for (size_t i = 0; i < in.size(); i++) {
        if (in[i] != '&') { out += in[i]; continue; }
        size_t semi = in.find(';', i);
        if (semi == std::string::npos) { out += in[i]; continue; }
        const std::string ent = in.substr(i + 1, semi - i - 1);
        if (ent == "amp") out += '&';
        else if (ent == "lt") out += '<';
        else if (ent == "gt") out += '>';
        else if (ent == "quot") out += '"';
        else if (ent == "apos") out += '\'';
        else { out += in.substr(i, semi - i + 1); }
        i = semi;
    }
    return out;
}

// skips comments, processing instructions, doctype and text between elements
void skipMisc(Cursor& c) {
    for (;;) {
        while (!c.eof() && c.s[c.i] != '<') c.i++;
        if (c.eof()) return;
        if (c.startsWith("<!--")) c.skipUntil("-->");
        else if (c.startsWith("<?")) c.skipUntil("?>");
        else if (c.startsWith("<!")) c.skipUntil(">");
        else return;
    }
}

std::unique_ptr<XmlNode> parseElement(Cursor& c) {
    if (c.eof() || c.s[c.i] != '<') c.fail("expected '<'");
    c.i++;
    auto node = std::make_unique<XmlNode>();
    while (!c.eof() && isNameChar(c.s[c.i])) node->tag += c.s[c.i++];
    if (node->tag.empty()) c.fail("empty element name");
    for (;;) {
        c.skipSpace();
        if (c.eof()) c.fail("unterminated start tag <" + node->tag);
        if (c.startsWith("/>")) { c.i += 2; return node; }
        if (c.s[c.i] == '>') { c.i++; break; }
        std::string name;
        while (!c.eof() && isNameChar(c.s[c.i])) name += c.s[c.i++];
        if (name.empty()) c.fail("bad attribute in <" + node->tag);
        c.skipSpace();
        if (c.eof() || c.s[c.i] != '=') c.fail("attribute '" + name + "' without value");
        c.i++;
        c.skipSpace();
        if (c.eof() || (c.s[c.i] != '"' && c.s[c.i] != '\'')) c.fail("attribute '" + name + "' value not quoted");
        const char quote = c.s[c.i++];
        size_t end = c.s.find(quote, c.i);
        if (end == std::string::npos) c.fail("unterminated attribute value");
        node->attrs.emplace_back(name, decodeEntities(c.s.substr(c.i, end - c.i)));
        c.i = end + 1;
    }
    for (;;) {
        skipMisc(c);
        if (c.eof()) c.fail("missing </" + node->tag + ">");
        if (c.startsWith("</")) {
            c.i += 2;
            std::string name;
            while (!c.eof() && isNameChar(c.s[c.i])) name += c.s[c.i++];
            if (name != node->tag) c.fail("mismatched </" + name + ">, open element is <" + node->tag + ">");
            c.skipSpace();
            if (c.eof() || c.s[c.i] != '>') c.fail("bad end tag");
            c.i++;
            return node;
        }
        node->children.push_back(parseElement(c));
    }
}

}  // namespace

std::unique_ptr<XmlNode> parseXml(const std::string& text) {
    Cursor c(text);
    skipMisc(c);
    if (c.eof()) throw std::runtime_error("xml: no root element");
    return parseElement(c);
}

}  // namespace gsbhost
