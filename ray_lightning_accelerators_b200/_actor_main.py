"""Entry point of a worker actor process:  python -m ray_lightning_accelerators_b200._actor_main ADDR KEYHEX

Started by _ray.ActorHandle with the environment already final (CUDA_VISIBLE_DEVICES narrowed
to the actor's GPU BEFORE the interpreter, and therefore torch, starts).  Listens on a unix
socket; every connection is one client (the driver, or a handle that travelled to another
worker).  Calls are executed one at a time, like a Ray actor.
"""
import os
import sys
import threading
import traceback
from multiprocessing.connection import Listener


def _serve(conn, state, lock):
    import cloudpickle
    while True:
        try:
            msg = conn.recv()
        except (EOFError, OSError):
            return
        kind = msg[0]
        if kind == "__exit__":
            try:
                conn.send((msg[1], True, None))
            except Exception:
                pass
            os._exit(0)
        call_id = msg[1]
        try:
            with lock:
                if kind == "__init__":
                    _, _, cls_payload, init_payload, sys_path = msg
                    for p in sys_path:
                        if p not in sys.path:
                            sys.path.append(p)
                    cls = cloudpickle.loads(cls_payload)
                    args, kwargs = cloudpickle.loads(init_payload)
                    state["obj"] = cls(*args, **kwargs)
                    out = None
                else:
                    _, _, name, payload = msg
                    args, kwargs = cloudpickle.loads(payload)
                    out = getattr(state["obj"], name)(*args, **kwargs)
            conn.send((call_id, True, cloudpickle.dumps(out)))
        except BaseException:
            try:
                conn.send((call_id, False, traceback.format_exc()))
            except Exception:
                return


def main():
    address, key = sys.argv[1], bytes.fromhex(sys.argv[2])
    listener = Listener(address, family="AF_UNIX", authkey=key)
    state, lock = {}, threading.Lock()
    while True:
        try:
            conn = listener.accept()
        except Exception:
            continue
        threading.Thread(target=_serve, args=(conn, state, lock), daemon=True).start()


if __name__ == "__main__":
    main()
