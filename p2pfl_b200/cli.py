"""Command line interface (reference: p2pfl/cli.py:71-202).

    python -m p2pfl_b200 experiment list
    python -m p2pfl_b200 experiment run mnist --nodes 4 --rounds 2 --epochs 1

``login`` / ``remote`` / ``launch`` are placeholders exactly as in the reference
(they print "not implemented").  ``experiment run`` starts the example in a
child interpreter and streams its output; unknown options are forwarded to it.
"""

from __future__ import annotations

import ast
import os
import subprocess
import sys
from typing import Dict, List

import typer
from rich.console import Console
from rich.table import Table

EXAMPLES_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "examples")
console = Console()

app = typer.Typer(
    help="p2pfl_b200 - decentralized federated learning on Blackwell GPUs",
    no_args_is_help=True,
    rich_markup_mode="markdown",
)
exp_app = typer.Typer(help="Run the bundled experiments.", no_args_is_help=True)
app.add_typer(exp_app, name="experiment")


def _not_implemented(what: str) -> None:
    console.print(f":warning: [yellow]{what} is not implemented yet[/yellow]")


@app.command()
def login(token: str = typer.Option("", help="API token of the web platform")) -> None:
    """Authenticate against the p2pfl web platform (placeholder)."""
    _not_implemented("login")


@app.command()
def remote() -> None:
    """Interact with a remote node (placeholder)."""
    _not_implemented("remote")


@app.command()
def launch(file: str = typer.Argument("", help="experiment description")) -> None:
    """Launch an experiment described in a file (placeholder)."""
    _not_implemented("launch")


def available_examples() -> Dict[str, str]:
    """example name -> first line of its module docstring."""
    out: Dict[str, str] = {}
    for fname in sorted(os.listdir(EXAMPLES_DIR)):
        if not fname.endswith(".py") or fname.startswith("_"):
            continue
        path = os.path.join(EXAMPLES_DIR, fname)
        try:
            doc = ast.get_docstring(ast.parse(open(path).read())) or ""
        except SyntaxError:
            doc = ""
        out[fname[:-3]] = doc.strip().splitlines()[0] if doc.strip() else ""
    return out


@exp_app.command("list")
def list_examples() -> None:
    """List the available examples."""
    table = Table("Name", "Description", title="examples")
    for name, desc in available_examples().items():
        table.add_row(name, desc)
    console.print(table)


@exp_app.command(
    "run", context_settings={"allow_extra_args": True, "ignore_unknown_options": True}
)
def run_example(ctx: typer.Context, example: str = typer.Argument(..., help="example name (see `experiment list`)")) -> None:
    """Run an example; extra options are passed through to it."""
    if example not in available_examples():
        console.print(f":x: [red]unknown example '{example}'[/red]")
        raise typer.Exit(code=1)
    cmd: List[str] = [sys.executable, os.path.join(EXAMPLES_DIR, example + ".py"), *ctx.args]
    console.print(f":rocket: running [bold]{' '.join(cmd)}[/bold]")
    env = dict(os.environ)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
    proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    assert proc.stdout is not None
    for line in proc.stdout:
        sys.stdout.write(line)
    code = proc.wait()
    if code != 0:
        console.print(f":x: [red]example exited with code {code}[/red]")
        raise typer.Exit(code=code)


def main() -> None:
    app()


if __name__ == "__main__":
    main()
