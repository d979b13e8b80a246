"""Support code: protobuf wire codec, event files, host-library binding, flags, timing."""
