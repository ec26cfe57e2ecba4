#include "xml_lite.hpp"

#include <cstring>
#include <stdexcept>

namespace tfk {
namespace xml {

const std::string *Element::attr(const char *name) const {
  for (const auto &kv : attrs)
    if (kv.first == name) return &kv.second;
  return nullptr;
}

const Element *Element::child(const char *tag_name) const {
  for (const auto &c : children)
    if (c->tag == tag_name) return c.get();
  return nullptr;
}

std::vector<const Element *> Element::children_named(const char *tag_name) const {
  std::vector<const Element *> out;
  for (const auto &c : children)
    if (c->tag == tag_name) out.push_back(c.get());
  return out;
}

namespace {

class Reader {
public:
  explicit Reader(const std::string &s) : s_(s), i_(0) {}

  std::unique_ptr<Element> document() {
    skip_misc();
    if (eof() || peek() != '<') fail("no root element");
    std::unique_ptr<Element> root = element();
    skip_misc();
    if (!eof()) fail("content after the root element");
    return root;
  }

private:
  const std::string &s_;
  size_t i_;

  bool eof() const { return i_ >= s_.size(); }
  char peek() const { return s_[i_]; }
  bool starts(const char *lit) const { return s_.compare(i_, std::strlen(lit), lit) == 0; }

  [[noreturn]] void fail(const std::string &what) const {
    size_t line = 1;
    for (size_t k = 0; k < i_ && k < s_.size(); ++k)
      if (s_[k] == '\n') ++line;
    throw std::invalid_argument("xml parse error (line " + std::to_string(line) + "): " + what);
  }

  static bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
  static bool is_name_char(char c) {
    return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == '_' || c == ':' ||
           c == '-' || c == '.' || (static_cast<unsigned char>(c) >= 0x80);
  }

  void skip_space() {
    while (!eof() && is_space(peek())) ++i_;
  }

  void skip_until(const char *close, const char *what) {
    size_t p = s_.find(close, i_);
    if (p == std::string::npos) fail(std::string("unterminated ") + what);
    i_ = p + std::strlen(close);
  }

  // whitespace, comments, processing instructions, DOCTYPE between top-level constructs
  void skip_misc() {
    for (;;) {
      skip_space();
      if (eof()) return;
      if (starts("<!--")) {
        i_ += 4;
        skip_until("-->", "comment");
      } else if (starts("<?")) {
        i_ += 2;
        skip_until("?>", "processing instruction");
      } else if (starts("<!DOCTYPE")) {
        // skip, honouring an internal subset in [...]
        int depth = 0;
        while (!eof()) {
          char c = s_[i_++];
          if (c == '[') ++depth;
          if (c == ']') --depth;
          if (c == '>' && depth <= 0) break;
        }
      } else {
        return;
      }
    }
  }

  std::string name() {
    size_t b = i_;
    while (!eof() && is_name_char(peek())) ++i_;
    if (i_ == b) fail("expected a name");
    return s_.substr(b, i_ - b);
  }

  static void append_utf8(std::string &out, unsigned long cp) {
    if (cp < 0x80) {
      out += static_cast<char>(cp);
    } else if (cp < 0x800) {
      out += static_cast<char>(0xC0 | (cp >> 6));
      out += static_cast<char>(0x80 | (cp & 0x3F));
    } else if (cp < 0x10000) {
      out += static_cast<char>(0xE0 | (cp >> 12));
      out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F));
      out += static_cast<char>(0x80 | (cp & 0x3F));
    } else {
      out += static_cast<char>(0xF0 | (cp >> 18));
      out += static_cast<char>(0x80 | ((cp >> 12) & 0x3F));
      out += static_cast<char>(0x80 | ((cp >> 6) & 0x3F));
      out += static_cast<char>(0x80 | (cp & 0x3F));
    }
  }

  std::string attr_value() {
    if (eof() || (peek() != '"' && peek() != '\'')) fail("expected a quoted attribute value");
    const char quote = s_[i_++];
    std::string out;
    while (!eof() && peek() != quote) {
      char c = s_[i_];
      if (c == '&') {
        size_t semi = s_.find(';', i_);
        if (semi == std::string::npos || semi - i_ > 10) {
          out += c; // lone ampersand: keep literally (TinyXML is equally lenient)
          ++i_;
          continue;
        }
        std::string ent = s_.substr(i_ + 1, semi - i_ - 1);
        if (ent == "lt")
          out += '<';
        else if (ent == "gt")
          out += '>';
        else if (ent == "amp")
          out += '&';
        else if (ent == "quot")
          out += '"';
        else if (ent == "apos")
          out += '\'';
        else if (!ent.empty() && ent[0] == '#') {
          unsigned long cp = (ent.size() > 1 && (ent[1] == 'x' || ent[1] == 'X'))
                                 ? std::strtoul(ent.c_str() + 2, nullptr, 16)
                                 : std::strtoul(ent.c_str() + 1, nullptr, 10);
          if (cp == 0 || cp > 0x10FFFF) fail("character reference &" + ent + "; is not a Unicode scalar value");
          append_utf8(out, cp);
        } else {
          out += s_.substr(i_, semi - i_ + 1); // unknown entity: keep literally
        }
        i_ = semi + 1;
      } else {
        out += c;
        ++i_;
      }
    }
    if (eof()) fail("unterminated attribute value");
    ++i_; // closing quote
    return out;
  }

  // URDF nests four or five levels deep; a bound keeps a crafted document from overflowing the stack (one frame per level)
  enum { MAX_DEPTH = 256 };
  int depth_ = 0;
  struct DepthGuard {
    int &d;
    explicit DepthGuard(int &x) : d(x) { ++d; }
    ~DepthGuard() { --d; }
  };

  std::unique_ptr<Element> element() {
    // precondition: peek() == '<' and this is a start tag
    DepthGuard guard(depth_);
    if (depth_ > MAX_DEPTH) fail("elements nested more than 256 levels deep");
    ++i_;
    std::unique_ptr<Element> e(new Element);
    e->tag = name();
    for (;;) {
      skip_space();
      if (eof()) fail("unterminated start tag <" + e->tag);
      if (starts("/>")) {
        i_ += 2;
        return e;
      }
      if (peek() == '>') {
        ++i_;
        break;
      }
      std::string key = name();
      skip_space();
      if (eof() || peek() != '=') fail("attribute '" + key + "' without '='");
      ++i_;
      skip_space();
      e->attrs.emplace_back(std::move(key), attr_value());
    }
    // content
    for (;;) {
      size_t lt = s_.find('<', i_);
      if (lt == std::string::npos) fail("unterminated element <" + e->tag + ">");
      i_ = lt; // text is ignored
      if (starts("<!--")) {
        i_ += 4;
        skip_until("-->", "comment");
      } else if (starts("<![CDATA[")) {
        i_ += 9;
        skip_until("]]>", "CDATA section");
      } else if (starts("<?")) {
        i_ += 2;
        skip_until("?>", "processing instruction");
      } else if (starts("</")) {
        i_ += 2;
        std::string close = name();
        if (close != e->tag) fail("mismatched end tag </" + close + "> for <" + e->tag + ">");
        skip_space();
        if (eof() || peek() != '>') fail("malformed end tag </" + close);
        ++i_;
        return e;
      } else {
        e->children.push_back(element());
      }
    }
  }
};

} // namespace

std::unique_ptr<Element> parse(const std::string &text) {
  Reader r(text);
  return r.document();
}

} // namespace xml
} // namespace tfk
