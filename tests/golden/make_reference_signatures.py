"""Extracts the public surface of the reference's module seam (class names, constructor parameters in order,
public methods of sevenn/nn/*.py; SevenNetCalculator ctor; deploy entry points) with `ast` -- no import of the
reference needed -- into tests/golden/reference_nn_signatures.json.  Authoring container only."""
import ast
import glob
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
REF = '/root/reference/sevenn'


def classes(path):
    out = {}
    tree = ast.parse(open(path).read())
    for n in tree.body:
        if isinstance(n, ast.ClassDef) and not n.name.startswith('_'):
            init = [m for m in n.body if isinstance(m, ast.FunctionDef) and m.name == '__init__']
            params, kwargs = [], False
            if init:
                a = init[0].args
                params = [x.arg for x in a.args][1:] + [x.arg for x in a.kwonlyargs]
                kwargs = a.kwarg is not None
            methods = sorted(m.name for m in n.body if isinstance(m, ast.FunctionDef) and not m.name.startswith('_'))
            out[n.name] = {'params': params, 'var_kwargs': kwargs, 'methods': methods}
    return out


def functions(path, names):
    tree = ast.parse(open(path).read())
    return {n.name: [x.arg for x in n.args.args] for n in tree.body
            if isinstance(n, ast.FunctionDef) and n.name in names}


def main():
    sig = {'nn': {}}
    for f in sorted(glob.glob(f'{REF}/nn/*.py')):
        if os.path.basename(f) in ('cue_helper.py', 'activation.py', 'util.py', '__init__.py'):
            continue
        sig['nn'].update(classes(f))
    sig['calculator'] = classes(f'{REF}/sevennet_calculator.py')['SevenNetCalculator']
    sig['deploy'] = functions(f'{REF}/scripts/deploy.py', ('deploy', 'deploy_parallel'))
    sig['model_build'] = functions(f'{REF}/model_build.py', ('build_E3_equivariant_model',))
    json.dump(sig, open(f'{HERE}/reference_nn_signatures.json', 'w'), indent=1, sort_keys=True)
    print({k: len(v) for k, v in sig.items()})


if __name__ == '__main__':
    main()
