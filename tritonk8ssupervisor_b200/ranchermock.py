"""A local stand-in for the five Rancher REST calls the playbook makes
(SURVEY.md §8f rank 1).

The roles talk to Rancher 1.x on ``http://<master>:8080`` with fixed shapes
(ansible/roles/ranchermaster/tasks/main.yml:29-52, roles/rancherhost/tasks/main.yml:6-34):

  1. ``GET  /v2-beta/projectTemplates?name=kubernetes``      -> ``.data[0].id``
  2. ``POST /v2-beta/projects`` (JSON body, expects **201**)  -> ``.id``
  3. ``POST /v1/registrationtokens?projectId=<id>`` (**201**) -> ``.links.self``
  4. ``GET  <links.self>``                                    -> ``.registrationUrl``
  5. the agent container is started with that URL as its command

Rancher is third-party and unpinned (``rancher/server:stable``), so the mock
restates only what the roles rely on: paths, methods, status codes and the JSON
members they read.  It validates what it is sent (members of the project body,
the ``projectId`` of the token request) and records every request, so a dry run
can assert the wire format.
"""
from __future__ import annotations

import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Any, Dict, List, Optional, Tuple
from urllib.parse import parse_qs, urlsplit

PROJECT_BODY_MEMBERS = {"description", "name", "projectTemplateId", "allowSystemRole", "members", "virtualMachine",
                        "servicesPortRange", "projectLinks"}


class RancherMock:
    """Threaded HTTP server on 127.0.0.1:<ephemeral>; use as a context manager."""

    def __init__(self, public_host: str = "10.0.0.1", public_port: int = 8080, template_id: str = "1pt5",
                 first_project: int = 0x1a7, fail: Optional[Dict[Tuple[str, str], int]] = None):
        #: how the roles address the server; links in responses use this authority
        self.public = "%s:%d" % (public_host, public_port)
        self.template_id = template_id
        self._next_project = first_project
        self._next_token = 1
        self.projects: Dict[str, Dict[str, Any]] = {}
        self.tokens: Dict[str, str] = {}          # token id -> project id
        self.requests: List[Dict[str, Any]] = []
        #: (method, path) -> status to answer instead (fault injection)
        self.fail = dict(fail or {})
        self._lock = threading.Lock()
        mock = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"

            def log_message(self, *a):  # silence
                pass

            def _send(self, status: int, doc: Any) -> None:
                body = json.dumps(doc).encode("utf-8")
                self.send_response(status)
                self.send_header("Content-Type", "application/json")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def _handle(self, method: str) -> None:
                n = int(self.headers.get("Content-Length") or 0)
                raw = self.rfile.read(n) if n else b""
                parts = urlsplit(self.path)
                status, doc = mock._dispatch(method, parts.path, parse_qs(parts.query), raw, dict(self.headers))
                self._send(status, doc)

            def do_GET(self):
                self._handle("GET")

            def do_POST(self):
                self._handle("POST")

        self._server = ThreadingHTTPServer(("127.0.0.1", 0), Handler)
        self.address = "127.0.0.1:%d" % self._server.server_address[1]
        self._thread = threading.Thread(target=self._server.serve_forever, daemon=True)

    # -- lifecycle ---------------------------------------------------------
    def __enter__(self) -> "RancherMock":
        self._thread.start()
        return self

    def __exit__(self, *exc) -> None:
        self._server.shutdown()
        self._server.server_close()

    # -- behaviour ---------------------------------------------------------
    def _dispatch(self, method: str, path: str, query: Dict[str, List[str]], raw: bytes,
                  headers: Dict[str, str]) -> Tuple[int, Any]:
        with self._lock:
            rec: Dict[str, Any] = {"method": method, "path": path, "query": {k: v[0] for k, v in query.items()},
                                   "headers": {k.lower(): v for k, v in headers.items()}, "body": raw.decode("utf-8", "replace")}
            self.requests.append(rec)
            if (method, path) in self.fail:
                return self.fail[(method, path)], {"type": "error", "status": self.fail[(method, path)]}
            if method == "GET" and path == "/v2-beta/projectTemplates":
                data = [{"id": self.template_id, "type": "projectTemplate", "name": "kubernetes"}]
                if query.get("name", ["kubernetes"])[0] != "kubernetes":
                    data = []
                return 200, {"type": "collection", "resourceType": "projectTemplate", "data": data}
            if method == "POST" and path == "/v2-beta/projects":
                try:
                    body = json.loads(raw.decode("utf-8"))
                except ValueError:
                    return 422, {"type": "error", "code": "InvalidBodyContent"}
                if not isinstance(body, dict) or not PROJECT_BODY_MEMBERS <= set(body):
                    return 422, {"type": "error", "code": "MissingRequired",
                                 "fieldName": sorted(PROJECT_BODY_MEMBERS - set(body if isinstance(body, dict) else []))}
                if body["projectTemplateId"] != self.template_id:
                    return 422, {"type": "error", "code": "InvalidReference", "fieldName": "projectTemplateId"}
                pid = "%x" % self._next_project
                self._next_project += 1
                self.projects[pid] = body
                rec["json"] = body
                return 201, {"id": pid, "type": "project", "name": body["name"], "description": body["description"],
                             "state": "registering"}
            if method == "POST" and path == "/v1/registrationtokens":
                pid = query.get("projectId", [""])[0]
                if pid not in self.projects:
                    return 404, {"type": "error", "code": "NotFound", "fieldName": "projectId"}
                tid = "1c%d" % self._next_token
                self._next_token += 1
                self.tokens[tid] = pid
                return 201, {"id": tid, "type": "registrationToken", "state": "registering",
                             "links": {"self": "http://%s/v1/registrationtokens/%s" % (self.public, tid)}}
            if method == "GET" and path.startswith("/v1/registrationtokens/"):
                tid = path.rsplit("/", 1)[1]
                if tid not in self.tokens:
                    return 404, {"type": "error", "code": "NotFound"}
                return 200, {"id": tid, "type": "registrationToken", "state": "active",
                             "registrationUrl": "http://%s/v1/scripts/TOKEN%s:%s" % (self.public, tid.upper(), self.tokens[tid]),
                             "links": {"self": "http://%s/v1/registrationtokens/%s" % (self.public, tid)}}
            return 404, {"type": "error", "code": "NotFound", "path": path}

    def wire_log(self) -> List[str]:
        """``METHOD path?query`` per request, in order."""
        out = []
        for r in self.requests:
            q = "&".join("%s=%s" % kv for kv in r["query"].items())
            out.append("%s %s%s" % (r["method"], r["path"], "?" + q if q else ""))
        return out
