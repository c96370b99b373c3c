// Test scaffolding for the oracle build only: minimal stand-ins for the protobuf / libvgio
// headers that rpvg's hot-path translation units include transitively. Not product code.
#pragma once
#include <istream>
namespace vg { namespace io {
template<class T> struct ProtobufIterator { ProtobufIterator(std::istream &) {} bool has_current() const { return false; } const T & operator*() const { return t; } ProtobufIterator & operator++() { return *this; } T t; };
}}
