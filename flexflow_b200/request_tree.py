"""Host-side SpecInfer tree logic of the reference's RequestManager, through the C ABI (include/ffb200.h):
initBitMask / updateBitMask / appendPendingRequest / appendBitMask (src/runtime/request_manager.cc:2382-2517) and
traverse_verify_tree (:2583-2741).  Integer only, runs on the host; the masks feed causalMask[] of the batch the
tree-verify / beam attention kernels read."""
import ctypes as C

from . import batch_config as bcfg
from . import lib as _lib


class TreeNode(C.Structure):
    _fields_ = [("token_id", C.c_int), ("depth", C.c_int)]


class CommitEntry(C.Structure):
    _fields_ = [("depth", C.c_int), ("index_in_batch", C.c_int)]


def initBitMask(bitmask: bcfg.BitMask, initLength: int):
    _lib.load().ffb_bitmask_init(C.addressof(bitmask), initLength)


def updateBitMask(bitmask: bcfg.BitMask, initLength: int, non_tree_size: int):
    _lib.load().ffb_bitmask_update(C.addressof(bitmask), initLength, non_tree_size)


def appendPendingRequest(bitmask: bcfg.BitMask, initLength: int):
    _lib.load().ffb_bitmask_append_pending(C.addressof(bitmask), initLength)


def appendBitMask(bitmask: bcfg.BitMask, newNodes: int, layer_nodes, currentDepth: int):
    """layer_nodes[i] = tree.treeLayers[i].nodes_num_this_layer for i < currentDepth."""
    arr = (C.c_int * max(1, len(layer_nodes)))(*layer_nodes)
    _lib.check(_lib.load().ffb_bitmask_append(C.addressof(bitmask), newNodes, arr, currentDepth), "ffb_bitmask_append")


def traverse_verify_tree(inputSerializedTree, outputSerializedTree, committed_tokens):
    """inputSerializedTree / outputSerializedTree: [(token_id, depth)] in layer order; committed_tokens:
    [(depth, index_in_batch)] per input node.  Returns (verifiedTree [(token_id, depth)], new committed_tokens)."""
    n_in, n_out = len(inputSerializedTree), len(outputSerializedTree)
    tin = (TreeNode * max(1, n_in))(*[TreeNode(t, d) for t, d in inputSerializedTree])
    tout = (TreeNode * max(1, n_out))(*[TreeNode(t, d) for t, d in outputSerializedTree])
    cin = (CommitEntry * max(1, len(committed_tokens)))(*[CommitEntry(d, i) for d, i in committed_tokens])
    ver = (TreeNode * (n_in + 1))()
    cout = (CommitEntry * (n_in + 1))()
    nv, nc = C.c_int(0), C.c_int(0)
    _lib.check(_lib.load().ffb_traverse_verify_tree(tin, n_in, tout, n_out, cin, ver, C.byref(nv), cout, C.byref(nc)),
               "ffb_traverse_verify_tree")
    return ([(ver[i].token_id, ver[i].depth) for i in range(nv.value)],
            [(cout[i].depth, cout[i].index_in_batch) for i in range(nc.value)])
