// Test scaffolding for the oracle build only: minimal stand-ins for the protobuf / libvgio
// headers that rpvg's hot-path translation units include transitively. Not product code.
#pragma once
#include <string>
namespace google { namespace protobuf {
class Message { public: virtual ~Message() {} virtual std::string GetTypeName() const { return "shim"; } };
template<class T> class RepeatedPtrField;
namespace util {
struct Status { bool ok() const { return false; } std::string ToString() const { return "shim"; } };
inline Status MessageToJsonString(const Message &, std::string *) { return Status(); }
inline Status JsonStringToMessage(const std::string &, Message *) { return Status(); }
}}}
