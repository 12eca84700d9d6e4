#!/bin/sh
# Builds the Java facade where a JDK exists (this image has none: no java/javac/jni.h).
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
: "${JAVA_HOME:?set JAVA_HOME to a JDK}"
python "$HERE/../build.py"
mkdir -p "$HERE/build/classes"
find "$HERE/src/main/java" -name '*.java' | xargs "$JAVA_HOME/bin/javac" -d "$HERE/build/classes"
g++ -O2 -fPIC -shared -std=c++17 -I"$JAVA_HOME/include" -I"$JAVA_HOME/include/linux" \
    "$HERE/jni/mp4j_jni.cc" -L"$HERE/../ytk_mp4j_b200/_lib" -lmp4j -Wl,-rpath,"$HERE/../ytk_mp4j_b200/_lib" \
    -o "$HERE/build/libmp4j_jni.so"
(cd "$HERE/build/classes" && "$JAVA_HOME/bin/jar" cf ../mp4j-b200.jar com)
echo "built $HERE/build/mp4j-b200.jar and libmp4j_jni.so (run with -Djava.library.path=$HERE/build)"
