"""Stand-in for the ``parameterized`` package the reference's test_neural_net_model.py imports (not installed
here and there is no network): ``@parameterized.expand(cases)`` on a unittest.TestCase method defines one test
method per case in the class body and unbinds the original name, as the real package does."""
import inspect


class parameterized:  # noqa: N801  (the package's own spelling)
    @staticmethod
    def expand(cases):
        cases = list(cases)

        def decorate(fn):
            class_body = inspect.currentframe().f_back.f_locals
            for i, case in enumerate(cases):
                args = tuple(case) if isinstance(case, (tuple, list)) else (case,)

                def make(bound):
                    def test(self):
                        return fn(self, *bound)
                    return test
                test = make(args)
                test.__name__ = f"{fn.__name__}_{i}"
                test.__doc__ = fn.__doc__
                class_body[test.__name__] = test
            return None
        return decorate
