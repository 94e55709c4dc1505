"""Headless command line (python -m synerfgine_b200 ≙ the reference's src/main.cu): argument handling without a GPU."""
from synerfgine_b200.__main__ import build_parser, main


def test_reference_flag_names_parse():
    a = build_parser().parse_args(["--load_snapshot", "a.ingp", "--rt", "scene.json", "--width", "640", "--height", "360", "--sshadows", "2", "--nshadows", "4"])
    assert (a.snapshot, a.virtual, a.width, a.height, a.sshadows, a.nshadows) == ("a.ingp", "scene.json", 640, 360, 2, 4)
    a = build_parser().parse_args(["--snapshot", "b.msgpack", "--virtual", "s.json"])
    assert (a.snapshot, a.virtual, a.width, a.height) == ("b.msgpack", "s.json", 1280, 720)          # main.cu defaults


def test_version_and_refusals(capsys):
    assert main(["--version"]) == 0 and "synerfgine_b200" in capsys.readouterr().out
    assert main([]) == 2 and "nothing to render" in capsys.readouterr().err
    assert main(["--train", "--synthetic", "small16x2"]) == 2 and "outside this library's scope" in capsys.readouterr().err
    assert main(["mesh.obj"]) == 2 and "only snapshots" in capsys.readouterr().err
