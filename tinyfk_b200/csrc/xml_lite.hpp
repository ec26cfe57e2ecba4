// xml_lite.hpp -- a small non-validating XML DOM reader, enough for URDF files.
//
// Host-side, one-time work (the reference vendors TinyXML for this;
// it is not on the hot path and not an acceleration target).  Supports: prolog / processing
// instructions, comments, DOCTYPE (skipped), CDATA and text (skipped: URDF carries all data in
// attributes), elements with single- or double-quoted attributes, the five predefined entities and
// numeric character references in attribute values.  Malformed input throws std::invalid_argument
// (the reference silently returns a null model and then dereferences it: src/tinyfk.cpp:20-27).
#ifndef TFK_XML_LITE_HPP
#define TFK_XML_LITE_HPP

#include <memory>
#include <string>
#include <utility>
#include <vector>

namespace tfk {
namespace xml {

struct Element {
  std::string tag;
  std::vector<std::pair<std::string, std::string>> attrs;
  std::vector<std::unique_ptr<Element>> children;

  // first attribute with that name, or nullptr
  const std::string *attr(const char *name) const;
  // first child element with that tag, or nullptr
  const Element *child(const char *tag_name) const;
  // all direct children with that tag, in document order
  std::vector<const Element *> children_named(const char *tag_name) const;
};

// Parses a whole document and returns its root element.
std::unique_ptr<Element> parse(const std::string &text);

} // namespace xml
} // namespace tfk
#endif
