ssh test@127.0.0.1 "kill -9 2666"
ssh test@127.0.0.1 "kill -9 2667"
ssh test@127.0.0.1 "kill -9 2668"
kill -9 2109
