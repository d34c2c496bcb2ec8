"""Same star-export surface as the reference's dino/construction/__init__.py:19-30."""
from .dataUtilities import *
from .jacobianConstruction import *
from .jacobianTrainingDriver import *
from .neuralNetworks import *
from .neuralNetworkUtilities import *
from .observableConstruction import *
from .trainingUtilities import *
