// stand-in for extern/owl/owl/include/owl/owl_device_buffer.h:24-32 (the real header drags in
// owl_host.h -> <optix.h>, absent here).  Same three fields in the same order.
#pragma once
#include <stddef.h>
typedef int OWLDataType;
namespace owl {
namespace device {
struct Buffer {
    OWLDataType type;
    size_t count;
    void* data;
};
}  // namespace device
}  // namespace owl
