#!/bin/sh
# Builds the runtime image and copies the wheel it produced to ./dist (reference: docker/build_images.sh, build_docker.sh).
set -e
cd "$(dirname "$0")/.."
TAG="${1:-metisfl_b200:latest}"
docker build -f docker/Dockerfile -t "$TAG" .
mkdir -p dist
CID=$(docker create "$TAG")
docker cp "$CID":/wheels/. dist/
docker rm "$CID" >/dev/null
echo "image $TAG built; wheel(s) in dist/"
