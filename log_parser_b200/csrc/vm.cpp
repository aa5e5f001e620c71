// Backtracking bytecode compiler + host interpreter (see vm.h).
#include "vm.h"

namespace lpx {

void build_vm(const ParsedRegex&, VmProgram&) {
    throw RegexUnsupported("pattern needs the backtracking verifier (look-around, back-reference, possessive/atomic, "
                           "MULTILINE anchors or an oversized DFA), which this build does not include yet");
}
int vm_match_host(const VmProgram&, const uint8_t*, size_t) { return -1; }

}  // namespace lpx
