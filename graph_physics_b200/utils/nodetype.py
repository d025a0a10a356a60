"""Node-type codes of the reference datasets (graphphysics/utils/nodetype.py:4-12); `SIZE` is the one-hot width."""
import enum


class NodeType(enum.IntEnum):
    NORMAL = 0
    OBSTACLE = 1
    AIRFOIL = 2
    HANDLE = 3
    INFLOW = 4
    OUTFLOW = 5
    WALL_BOUNDARY = 6
    SIZE = 9
