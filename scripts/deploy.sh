#!/bin/bash
# Redeploy the LEARN demo container on a host (counterpart of the reference's scripts/deploy.sh): pull the image, replace the
# running container. usage: scripts/deploy.sh <user@host> <image> ; the ssh key comes from $DEPLOY_KEY (private key text) if set.
set -euo pipefail
HOST=${1:?user@host}
IMAGE=${2:?image}
KEYFILE=""
if [ -n "${DEPLOY_KEY:-}" ]; then
  KEYFILE=$(mktemp)
  trap 'rm -f "$KEYFILE"' EXIT
  printf '%s\n' "$DEPLOY_KEY" > "$KEYFILE"
  chmod 600 "$KEYFILE"
fi
ssh ${KEYFILE:+-i "$KEYFILE"} -o StrictHostKeyChecking=accept-new "$HOST" "
  docker pull $IMAGE &&
  (docker rm -f garfield-demo 2>/dev/null || true) &&
  docker run -d --restart unless-stopped --name garfield-demo -p 80:8080 $IMAGE"
