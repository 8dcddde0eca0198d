import importlib.util, sys, os
def load():
    if 'multibody_jl_b200' in sys.modules: return sys.modules['multibody_jl_b200']
    root='/root/repo/multibody.jl_b200'
    spec=importlib.util.spec_from_file_location('multibody_jl_b200', root+'/__init__.py', submodule_search_locations=[root])
    m=importlib.util.module_from_spec(spec); sys.modules['multibody_jl_b200']=m; spec.loader.exec_module(m); return m
