"""Load the reference's protobuf text configs on a modern protobuf.

The reference's generated ``second/protos/*_pb2.py`` are from a pre-3.19 protoc
and do not import under protobuf >= 4.  Each of them embeds its
FileDescriptorProto as a ``serialized_pb=_b('...')`` literal; this module lifts
that literal out with ``ast`` (nothing from the reference is executed), rebuilds
the descriptors in a private pool and returns real message classes, so
``text_format.Merge(config_text, TrainEvalPipelineConfig())`` works again and the
reference's ``.config`` files drive the new voxelizer unchanged.
"""
import ast
import glob
import os


def _serialized_pb(path):
    tree = ast.parse(open(path, "r", encoding="utf-8").read())
    for node in ast.walk(tree):
        if isinstance(node, ast.keyword) and node.arg == "serialized_pb":
            v = node.value
            if isinstance(v, ast.Call) and v.args:      # _b('...')
                v = v.args[0]
            s = ast.literal_eval(v)
            return s.encode("latin1") if isinstance(s, str) else s
    raise ValueError(f"no serialized_pb literal in {path}")


def load_pool(second_root):
    """DescriptorPool holding every message of ``<second_root>/second/protos``."""
    from google.protobuf import descriptor_pb2, descriptor_pool
    files = {}
    for p in sorted(glob.glob(os.path.join(second_root, "second", "protos", "*_pb2.py"))):
        fd = descriptor_pb2.FileDescriptorProto()
        fd.ParseFromString(_serialized_pb(p))
        files[fd.name] = fd
    pool = descriptor_pool.DescriptorPool()
    done = set()

    def add(name):
        if name in done or name not in files:
            return
        for dep in files[name].dependency:
            add(dep)
        pool.Add(files[name])
        done.add(name)

    for name in files:
        add(name)
    return pool


def message_class(pool, full_name):
    from google.protobuf import message_factory
    return message_factory.GetMessageClass(pool.FindMessageTypeByName(full_name))


def load_pipeline_config(config_path, second_root):
    """Parse a reference ``.config`` file into a TrainEvalPipelineConfig message."""
    from google.protobuf import text_format
    pool = load_pool(second_root)
    cls = message_class(pool, "second.protos.TrainEvalPipelineConfig")
    msg = cls()
    text_format.Merge(open(config_path, "r").read(), msg)
    return msg
