#!/usr/bin/env bash
# Same command line as the reference's run.sh (run.sh:2): input.gltf width height samples [output.ppm]
# build/solution is the C++ host in front of librtb.so (raytracing-hw_b200/host/main.cpp); ./build.sh makes both.
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
exec "$here/build/solution" "$@"
