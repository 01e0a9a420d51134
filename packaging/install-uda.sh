#!/usr/bin/env bash
# Installs an unpacked uda_b200 tree and wires it into a Hadoop installation (counterpart of the reference's
# build/install-uda.sh: option parsing, default directories, and — its lines 135-142 — an `uda-env.sh` that puts the
# plug-in jar on HADOOP_CLASSPATH plus one guarded `source` line appended to hadoop-env.sh, added only once).
#
#   install-uda.sh --uda-dir=DIR [--prefix=/] [--uda-hadoop-conf-dir=DIR] [--lib-dir=DIR] [--doc-dir=DIR] [--build-dir=DIR]
#
#   --uda-dir              where the package payload lives / is copied to (libuda.so, libuda_kernels.so, uda_b200/, java/)
#   --build-dir            source tree to copy from (default: the directory above this script; skipped when it is --uda-dir)
#   --uda-hadoop-conf-dir  Hadoop configuration directory whose hadoop-env.sh gets the source line (optional)
set -euo pipefail

usage() { sed -n '2,12p' "$0"; exit 1; }

UDA_DIR="" PREFIX="/" HADOOP_CONF="" LIB_DIR="" DOC_DIR="" BUILD_DIR="$(cd "$(dirname "$0")/.." && pwd)"
for arg in "$@"; do
    case "$arg" in
        --uda-dir=*) UDA_DIR="${arg#*=}" ;;
        --prefix=*) PREFIX="${arg#*=}" ;;
        --uda-hadoop-conf-dir=*) HADOOP_CONF="${arg#*=}" ;;
        --lib-dir=*) LIB_DIR="${arg#*=}" ;;
        --doc-dir=*) DOC_DIR="${arg#*=}" ;;
        --build-dir=*) BUILD_DIR="${arg#*=}" ;;
        -h|--help) usage ;;
        *) echo "Unknown option: $arg" >&2; usage ;;
    esac
done
[ -n "$UDA_DIR" ] || { echo "Missing param: --uda-dir" >&2; usage; }
PREFIX="${PREFIX%/}"
LIB_DIR="${LIB_DIR:-$PREFIX/usr/lib}"
DOC_DIR="${DOC_DIR:-$PREFIX/usr/share/doc/uda-b200}"

mkdir -p "$UDA_DIR" "$DOC_DIR"
if [ "$(cd "$BUILD_DIR" && pwd)" != "$(cd "$UDA_DIR" && pwd)" ]; then
    # copy the payload: native libraries, Python package, plug-in sources / jars, utilities, docs
    for lib in libuda.so libuda_kernels.so; do
        [ -f "$BUILD_DIR/uda_b200/lib/$lib" ] && install -m 0755 "$BUILD_DIR/uda_b200/lib/$lib" "$UDA_DIR/$lib"
        [ -f "$BUILD_DIR/$lib" ] && install -m 0755 "$BUILD_DIR/$lib" "$UDA_DIR/$lib"
    done
    for d in uda_b200 java utils; do
        [ -d "$BUILD_DIR/$d" ] && { mkdir -p "$UDA_DIR/$d"; (cd "$BUILD_DIR/$d" && tar --exclude='__pycache__' --exclude='*.so' -cf - .) | (cd "$UDA_DIR/$d" && tar -xf -); }
    done
    for f in README.md INTEGRATION.md DESIGN.md; do
        [ -f "$BUILD_DIR/$f" ] && install -m 0644 "$BUILD_DIR/$f" "$DOC_DIR/$f"
    done
fi

# the environment file Hadoop's launch scripts source: class path of the plug-in and where the JVM finds libuda.so
ENV_FILE="$UDA_DIR/uda-env.sh"
{
    echo "# written by install-uda.sh"
    echo "export HADOOP_CLASSPATH=\$HADOOP_CLASSPATH:${UDA_DIR}/uda.jar"
    echo "export JAVA_LIBRARY_PATH=\${JAVA_LIBRARY_PATH:+\$JAVA_LIBRARY_PATH:}${UDA_DIR}"
    echo "export LD_LIBRARY_PATH=\${LD_LIBRARY_PATH:+\$LD_LIBRARY_PATH:}${UDA_DIR}"
    echo "export UDA_KERNELS_LIB=${UDA_DIR}/libuda_kernels.so"
} > "$ENV_FILE"

if [ -n "$HADOOP_CONF" ]; then
    HENV="$HADOOP_CONF/hadoop-env.sh"
    if [ -w "$HENV" ] && ! grep --quiet "uda-env.sh" "$HENV"; then
        printf '\nif [ -f %s ]; then . %s; fi\n' "$ENV_FILE" "$ENV_FILE" >> "$HENV"
        echo "hooked $ENV_FILE into $HENV"
    elif [ -w "$HENV" ]; then
        echo "$HENV already sources uda-env.sh"
    else
        echo "warning: $HENV is not writable; add '. $ENV_FILE' to it by hand" >&2
    fi
fi
echo "uda_b200 installed in $UDA_DIR"
